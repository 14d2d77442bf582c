// spgemm_numeric.cuh -- numeric Gustavson pass: C_i = sum_k a_ik * B_k with sorted columns.
//
// Takes the place of the value half of cusparseSpGEMM_compute and of cusparseSpGEMM_copy
// (reference: cupy_cusparse/spgemm_from_txt_alg1.cu:177-188, cupyx/cusparse.py:2108-2137).
//
// Ordering contract (see common.cuh): every output entry is the sequential sum, in A-row
// storage order, of separately rounded products.  That makes the values bit-identical to
// scipy.sparse (the reference's CPU path) and run-to-run deterministic: no floating-point
// atomics anywhere; parallelism comes from distinct columns only.
#pragma once
#include <limits.h>
#include "common.cuh"

namespace b200sp {

template <typename T>
__device__ __forceinline__ T scale_out(T alpha, T v) {
  return alpha == T(1) ? v : mul_rn(alpha, v);
}

// In-place stable compaction of the valid (key != EMPTY) slots of a group's table to the
// front; returns the number of valid entries.  Safe in place: the write position never
// exceeds the read position and each G-slot chunk is read in full before it is written.
template <int G, typename T>
__device__ __forceinline__ int compact_table(int *keys, T *vals, int S) {
  const int lane = lane_id();
  const int sub = lane & (G - 1);
  int n_valid = 0;
  for (int c = 0; c < S; c += G) {
    const int k = keys[c + sub];
    const T v = vals[c + sub];
    const bool valid = (k != EMPTY);
    const unsigned bal = __ballot_sync(FULL, valid);
    unsigned gb;
    if (G == 32) gb = bal;
    else gb = (bal >> (lane & ~(G - 1))) & ((1u << (G & 31)) - 1u);
    const int pos = n_valid + __popc(gb & ((1u << sub) - 1u));
    __syncwarp();
    if (valid) {
      keys[pos] = k;
      vals[pos] = v;
    }
    n_valid += __popc(gb);
    __syncwarp();
  }
  return n_valid;
}

// Bitonic sort of P = 2^x (key, val) pairs in shared memory by `nthr` threads (tid in
// [0,nthr)); SYNC() is __syncwarp or __syncthreads.
template <typename T, class SyncF>
__device__ __forceinline__ void bitonic_sort(int *keys, T *vals, int P, int tid, int nthr,
                                             SyncF &&sync) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = tid; t < (P >> 1); t += nthr) {
        const int idx = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int ixj = idx | j;
        const int ka = keys[idx], kb = keys[ixj];
        const bool up = (idx & k) == 0;
        if ((ka > kb) == up) {
          keys[idx] = kb;
          keys[ixj] = ka;
          const T va = vals[idx];
          vals[idx] = vals[ixj];
          vals[ixj] = va;
        }
      }
      sync();
    }
  }
}

// ---- group-of-G-lanes per row, hash accumulator of S slots per group ---------------------
// G = 8,  S = 64   : rows of C with <= 32 entries (4 rows per warp), rank sort
// G = 32, S = 256+ : rows of C with <= S/2 entries (warp per row), rank or bitonic sort
// RANKBM (G == 32, few columns): a per-warp bitmap over the n columns records the touched
// columns; the rank of a column among the set bits IS its position in the sorted output row, so
// the table is emitted straight to C without compaction or sort (the bitonic sort of a 512-entry
// row costs ~10x the accumulation itself).
template <int G, typename T, bool RANKBM>
__global__ void __launch_bounds__(128)
k_num_hash_group(const int *__restrict__ rows, int nrows, int S, int n, const int *__restrict__ Ap,
                 const int *__restrict__ Aj, const T *__restrict__ Ax,
                 const int *__restrict__ Bp, const int *__restrict__ Bj,
                 const T *__restrict__ Bx, const int *__restrict__ Cp, int *__restrict__ Cj,
                 T *__restrict__ Cx, T alpha) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int groups_per_block = 128 / G;
  const int g = threadIdx.x / G;
  const int lane = lane_id();
  const int sub = lane & (G - 1);
  T *vals = reinterpret_cast<T *>(smem_raw) + g * S;
  int *keys = reinterpret_cast<int *>(smem_raw + (size_t)groups_per_block * S * sizeof(T)) + g * S;
  const unsigned mask = S - 1;
  const int uid_base = (lane / G) * S;  // unique across the groups of this warp
  const int ridx = blockIdx.x * groups_per_block + g;
  const int row = ridx < nrows ? rows[ridx] : -1;
  for (int s = sub; s < S; s += G) {
    keys[s] = EMPTY;
    vals[s] = T(0);
  }
  const int nwords = (n + 31) >> 5;
  unsigned *bm = nullptr;
  int *pre = nullptr;
  if (RANKBM) {
    bm = reinterpret_cast<unsigned *>(smem_raw + (size_t)groups_per_block * S * (sizeof(T) + 4)) + (size_t)g * 2 * nwords;
    pre = reinterpret_cast<int *>(bm + nwords);
    for (int s = sub; s < nwords; s += G) bm[s] = 0u;
  }
  __syncwarp();
  int a_begin = 0, a_end = 0;
  if (row >= 0) {
    a_begin = __ldg(Ap + row);
    a_end = __ldg(Ap + row + 1);
  }
  struct Loaded {
    int j;
    T a, b;
  };
  stream_products<G>(
      a_begin, a_end, Aj, Bp,
      [&](bool act, int bidx, int aidx) {
        Loaded v;
        v.j = act ? __ldg(Bj + bidx) : 0;
        v.a = act ? __ldg(Ax + aidx) : T(0);
        v.b = act ? __ldg(Bx + bidx) : T(0);
        return v;
      },
      [&](bool act, const Loaded &ld, bool mc) {
        int slot = 0;
        T v = T(0);
        if (act) {
          bool fresh;
          slot = hash_insert(keys, mask, ld.j, fresh);
          if (RANKBM && fresh) atomicOr(&bm[ld.j >> 5], 1u << (ld.j & 31));
          v = mul_rn(ld.a, ld.b);
        }
        __syncwarp();  // inserts of this step are visible before the accumulate
        ordered_add(vals, slot, uid_base + slot, act, v, mc);
      });
  if (RANKBM) {
    if (row < 0) return;  // G == 32: warp-uniform
    const int out = __ldg(Cp + row);
    int carry = 0;  // exclusive prefix of the per-word popcounts
    for (int c = 0; c < nwords; c += 32) {
      const int wi = c + sub;
      const int cnt = wi < nwords ? __popc(bm[wi]) : 0;
      int incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(FULL, incl, o);
        if (sub >= o) incl += t;
      }
      if (wi < nwords) pre[wi] = carry + incl - cnt;
      carry += __shfl_sync(FULL, incl, 31);
    }
    __syncwarp();
    for (int s = sub; s < S; s += 32) {
      const int k = keys[s];
      if (k != EMPTY) {
        const int r = pre[k >> 5] + __popc(bm[k >> 5] & ((1u << (k & 31)) - 1u));
        Cj[out + r] = k;
        Cx[out + r] = scale_out(alpha, vals[s]);
      }
    }
    return;
  }
  const int nnz = compact_table<G, T>(keys, vals, S);
  if (row < 0) return;
  const int out = __ldg(Cp + row);
  if (G < 32 || nnz <= 64) {
    // rank sort: position = number of smaller keys (keys are distinct)
    for (int e = sub; e < nnz; e += G) {
      const int k = keys[e];
      int rank = 0;
      for (int f = 0; f < nnz; ++f) rank += (keys[f] < k);
      Cj[out + rank] = k;
      Cx[out + rank] = scale_out(alpha, vals[e]);
    }
  } else {
    const int P = next_pow2(nnz);
    for (int s = nnz + lane; s < P; s += 32) keys[s] = INT_MAX;
    __syncwarp();
    bitonic_sort(keys, vals, P, lane, 32, [] { __syncwarp(); });
    for (int e = lane; e < nnz; e += 32) {
      Cj[out + e] = keys[e];
      Cx[out + e] = scale_out(alpha, vals[e]);
    }
  }
}

// ---- CTA per row, hash accumulator of S slots in shared memory -----------------------------
// Rows of C with up to 0.75 * S entries.  The A nonzeros are walked one at a time (k-step):
// all threads stride the one B row, whose columns are distinct, so every accumulator is
// touched by at most one thread per step and the barrier between steps fixes the order.
template <typename T>
__global__ void __launch_bounds__(256)
k_num_hash_cta(const int *__restrict__ rows, int nrows, int S, const int *__restrict__ Ap,
               const int *__restrict__ Aj, const T *__restrict__ Ax, const int *__restrict__ Bp,
               const int *__restrict__ Bj, const T *__restrict__ Bx, const int *__restrict__ Cp,
               int *__restrict__ Cj, T *__restrict__ Cx, T alpha) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_start[256];
  __shared__ int s_len[256];
  __shared__ int s_nnz;
  T *vals = reinterpret_cast<T *>(smem_raw);
  int *keys = reinterpret_cast<int *>(smem_raw + (size_t)S * sizeof(T));
  const unsigned mask = S - 1;
  const int tid = threadIdx.x;
  for (int ridx = blockIdx.x; ridx < nrows; ridx += gridDim.x) {
    const int row = rows[ridx];
    for (int s = tid; s < S; s += 256) {
      keys[s] = EMPTY;
      vals[s] = T(0);
    }
    const int a_begin = __ldg(Ap + row), a_end = __ldg(Ap + row + 1);
    __syncthreads();
    for (int base = a_begin; base < a_end; base += 256) {
      const int t = base + tid;
      int st = 0, ln = 0;
      if (t < a_end) {
        const int kk = __ldg(Aj + t);
        st = __ldg(Bp + kk);
        ln = __ldg(Bp + kk + 1) - st;
      }
      s_start[tid] = st;
      s_len[tid] = ln;
      __syncthreads();
      const int cnt = min(256, a_end - base);
      for (int q = 0; q < cnt; ++q) {
        const int ql = s_len[q];
        if (ql == 0) continue;  // CTA-uniform
        const int qs = s_start[q];
        const T a = __ldg(Ax + base + q);
        for (int u = tid; u < ql; u += 256) {
          bool fresh;
          const int slot = hash_insert(keys, mask, __ldg(Bj + qs + u), fresh);
          vals[slot] = add_rn(vals[slot], mul_rn(a, __ldg(Bx + qs + u)));
        }
        __syncthreads();
      }
      __syncthreads();
    }
    // compact (warp 0, in place), then sort with the whole CTA
    if (tid < 32) {
      const int nn = compact_table<32, T>(keys, vals, S);
      if (tid == 0) s_nnz = nn;
    }
    __syncthreads();
    const int nnz = s_nnz;
    const int P = next_pow2(nnz);
    for (int s = nnz + tid; s < P; s += 256) keys[s] = INT_MAX;
    __syncthreads();
    bitonic_sort(keys, vals, P, tid, 256, [] { __syncthreads(); });
    const int out = __ldg(Cp + row);
    for (int e = tid; e < nnz; e += 256) {
      Cj[out + e] = keys[e];
      Cx[out + e] = scale_out(alpha, vals[e]);
    }
    __syncthreads();
  }
}

// ---- CTA per row, bitmap + rank, everything in global memory (any n, any row size) ---------
// Fallback for rows too large for a shared-memory table.  Per CTA workspace: nwords bitmap
// words + nwords exclusive popcount prefixes.  The rank of column j among the set bits is its
// position in the sorted output row, so values accumulate straight into C.data (no hash, no
// sort), ordered by the same k-step barrier as above.  The bitmap is zero on entry to each row.
template <typename T>
__global__ void __launch_bounds__(256)
k_num_bitmap_global(const int *__restrict__ rows, int nrows, int n, const int *__restrict__ Ap,
                    const int *__restrict__ Aj, const T *__restrict__ Ax,
                    const int *__restrict__ Bp, const int *__restrict__ Bj,
                    const T *__restrict__ Bx, const int *__restrict__ Cp, int *__restrict__ Cj,
                    T *Cx, T alpha, unsigned *__restrict__ gws) {
  __shared__ int s_start[256];
  __shared__ int s_len[256];
  __shared__ int s_scan[9];
  const int nwords = (n + 31) >> 5;
  unsigned *bm = gws + (size_t)blockIdx.x * 2 * nwords;
  int *prefix = reinterpret_cast<int *>(bm + nwords);
  const int tid = threadIdx.x, lane = lane_id(), w = tid >> 5;
  for (int ridx = blockIdx.x; ridx < nrows; ridx += gridDim.x) {
    const int row = rows[ridx];
    const int a_begin = __ldg(Ap + row), a_end = __ldg(Ap + row + 1);
    const int out = __ldg(Cp + row), nnz = __ldg(Cp + row + 1) - out;
    // (1) structure: set bits, warps take whole B rows
    for (int t = a_begin + w; t < a_end; t += 8) {
      const int kk = __ldg(Aj + t);
      const int be = __ldg(Bp + kk + 1);
      for (int u = __ldg(Bp + kk) + lane; u < be; u += 32) {
        const int j = __ldg(Bj + u);
        atomicOr(&bm[j >> 5], 1u << (j & 31));
      }
    }
    for (int e = tid; e < nnz; e += 256) Cx[out + e] = T(0);
    __syncthreads();
    // (2) exclusive prefix of the per-word popcounts (block scan with carry)
    int carry = 0;
    for (int s0 = 0; s0 < nwords; s0 += 256) {
      const int s = s0 + tid;
      const int c = s < nwords ? __popc(__ldcg(bm + s)) : 0;
      int incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += v;
      }
      if (lane == 31) s_scan[w] = incl;
      __syncthreads();
      if (tid == 0) {
        int run = 0;
        for (int q = 0; q < 8; ++q) {
          const int x = s_scan[q];
          s_scan[q] = run;
          run += x;
        }
        s_scan[8] = run;
      }
      __syncthreads();
      if (s < nwords) prefix[s] = carry + s_scan[w] + incl - c;
      carry += s_scan[8];
      __syncthreads();
    }
    // (3) ordered accumulation, one A nonzero per step
    for (int base = a_begin; base < a_end; base += 256) {
      const int t = base + tid;
      int st = 0, ln = 0;
      if (t < a_end) {
        const int kk = __ldg(Aj + t);
        st = __ldg(Bp + kk);
        ln = __ldg(Bp + kk + 1) - st;
      }
      s_start[tid] = st;
      s_len[tid] = ln;
      __syncthreads();
      const int cnt = min(256, a_end - base);
      for (int q = 0; q < cnt; ++q) {
        const int ql = s_len[q];
        if (ql == 0) continue;
        const int qs = s_start[q];
        const T a = __ldg(Ax + base + q);
        for (int u = tid; u < ql; u += 256) {
          const int j = __ldg(Bj + qs + u);
          const unsigned bits = __ldcg(bm + (j >> 5));
          const int r = __ldcg(prefix + (j >> 5)) + __popc(bits & ((1u << (j & 31)) - 1u));
          T *dst = Cx + out + r;
          __stcg(dst, add_rn(__ldcg(dst), mul_rn(a, __ldg(Bx + qs + u))));
        }
        __syncthreads();
      }
      __syncthreads();
    }
    // (4) emit sorted columns, scale, and clear the bitmap for the next row
    for (int s = tid; s < nwords; s += 256) {
      unsigned bits = __ldcg(bm + s);
      if (bits) {
        int p = out + __ldcg(prefix + s);
        bm[s] = 0u;
        while (bits) {
          const int b = __ffs(bits) - 1;
          bits &= bits - 1;
          Cj[p++] = (s << 5) + b;
        }
      }
    }
    if (alpha != T(1))
      for (int e = tid; e < nnz; e += 256) Cx[out + e] = mul_rn(alpha, __ldcg(Cx + out + e));
    __syncthreads();
  }
}

// ---- CTA per row, dense accumulator over the n columns in shared memory ---------------------
// For matrices whose column count fits shared memory (n * sizeof(T) <= ~200 KB) and rows of C
// with more than a few hundred entries: slot = column, so there is no hashing, no probing and
// no sort - the output sweep walks the columns in order.  The A nonzeros are applied one B-row
// piece at a time with a barrier in between (columns inside a B row are distinct), which keeps
// the per-column order of the sums = A-row storage order.
//
// An untouched column holds -0.0 and every product is canonicalised
// (p + 0.0 turns a -0.0 product into +0.0 and changes nothing else), so that
//   (-0.0) + p' = (+0.0) + p   bit for bit,
// the first add needs no special case, and an accumulator that has been touched can never be
// -0.0 again (x + y = -0.0 only for x = y = -0.0): explicit zeros from cancellation stay stored,
// no flag array or atomic is needed, and the read-modify-write chain is LDS -> FADD -> STS with no
// compare / select in it.
template <typename T> struct NegZero;
template <> struct NegZero<float> {
  static __device__ __forceinline__ float value() { return __int_as_float((int)0x80000000); }
  static __device__ __forceinline__ bool is(float v) { return __float_as_int(v) == (int)0x80000000; }
};
template <> struct NegZero<double> {
  static __device__ __forceinline__ double value() { return __longlong_as_double((long long)0x8000000000000000ULL); }
  static __device__ __forceinline__ bool is(double v) { return __double_as_longlong(v) == (long long)0x8000000000000000ULL; }
};

constexpr int DENSE_MAX_NT = 192;  // 3 CTAs of 6 warps: at most 5 warps per scheduler partition, 96 registers each
// B-row pieces whose loads are in flight ahead of the one being applied: DEPTH template parameter of the kernel, 8 or 12.
// Twelve when shared memory already limits the SM to three CTAs (n > 12 K columns in fp32: the extra registers are
// free), eight when more CTAs fit.  One call, variants built from the same sources (tools/gpu_r2_ac.sh): C3 d = 1e-2
// kernel 1.101 -> 1.057 ms with 12, C4 8192^2 d = 3e-2 0.915 -> 0.959 ms (four CTAs of 79 registers became three).
constexpr int DENSE_DEPTH_MAX = 12;
constexpr int DENSE_PAD = 2 * DENSE_DEPTH_MAX;  // zero-length staging entries after the last piece: the prefetch reads up to cnt + 2 D - 1
constexpr int DENSE_CAP = 2 * DENSE_MAX_NT;  // staged pieces per round

// Global loads that stay where they are written: the compiler sinks read-only (ld.global.nc)
// loads across bar.sync to shorten register lifetimes, which collapses a software pipeline that
// is built around barriers; a volatile asm load keeps its place in the instruction stream.
__device__ __forceinline__ int ld_inplace(const int *p) {
  int v;
  asm volatile("ld.global.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_inplace(const float *p) {
  float v;
  asm volatile("ld.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double ld_inplace(const double *p) {
  double v;
  asm volatile("ld.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}

// One staged piece of a B row (at most blockDim.x entries): where it starts, how long it is, and
// the A value it is scaled by (read back with one LDS.128).
template <typename T> struct __align__(16) DenseSeg {
  int start, len;
  T a;
};

// ---- packed, bank-dealt copy of B for k_num_dense_cta --------------------------------------------
// The L1 data pipe is what bounds the kernel (73 % busy on C3: 243 M wavefronts per launch, of them
// 78 M bank-conflict replays of the accumulator read-modify-write and ~65 M for the two 4-byte loads
// of a (column, value) pair, each straddling two 128-byte lines).  Both come from how a B row lies
// in memory, and the order of the entries INSIDE a B row changes no bit of C (its columns are
// distinct), so every numeric call first writes B once more (one kernel, ~2 x 8 nnzB bytes):
//   * (column, value) interleaved: one 8-byte (fp64: 16-byte) load per product instead of two;
//   * every row starts on a 32-entry boundary and is padded to whole groups of 32 entries (column -1
//     = no entry), so a warp's load is one aligned 256-byte span;
//   * inside a row the entries are dealt over its groups of 32 like cards, sorted by accumulator
//     bank (column mod 32; fp64: mod 16): a group - what one warp applies in one step - holds every
//     bank about once.
// Row kk occupies [ru32(Bp[kk]) + 32 kk, + ru32(len)) in entries: closed form, no scan, no second
// pointer array; the 32 spare entries per row are never read.
template <typename T> struct PBEnt;
template <> struct __align__(8) PBEnt<float> {
  int col;
  float val;
};
template <> struct __align__(16) PBEnt<double> {
  int col, pad;
  double val;
};
__device__ __forceinline__ int pb_start(int bp, int kk) { return ((bp + 31) & ~31) + 32 * kk; }
__device__ __forceinline__ void ld_pb(const PBEnt<float> *p, int &col, float &val) {
  asm volatile("ld.global.v2.b32 {%0, %1}, [%2];" : "=r"(col), "=f"(val) : "l"(p));
}
__device__ __forceinline__ void ld_pb(const PBEnt<double> *p, int &col, double &val) {
  long long c;
  asm volatile("ld.global.v2.b64 {%0, %1}, [%2];" : "=l"(c), "=d"(val) : "l"(p));
  col = (int)c;
}

template <typename T>
__global__ void __launch_bounds__(256)
k_pack_b(int k, const int *__restrict__ Bp, const int *__restrict__ Bj, const T *__restrict__ Bx,
         PBEnt<T> *__restrict__ pb) {
  constexpr int BANKS = sizeof(T) == 4 ? 32 : 16;
  __shared__ int s_base[8][32];
  const int lane = lane_id(), w = threadIdx.x >> 5;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const unsigned lt = (1u << lane) - 1u;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < k; r += warps) {
    const int b = __ldg(Bp + r), L = __ldg(Bp + r + 1) - b;
    if (L <= 0) continue;
    PBEnt<T> *row = pb + pb_start(b, r);
    const int G = (L + 31) >> 5;  // groups of 32 entries
    const float rG = 1.0f / (float)G;
    s_base[w][lane] = 0;
    __syncwarp();
    for (int t = lane; t < L; t += 32) atomicAdd(&s_base[w][__ldg(Bj + b + t) & (BANKS - 1)], 1);
    __syncwarp();
    const int c = s_base[w][lane];
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(FULL, inc, o);
      if (lane >= o) inc += u;
    }
    __syncwarp();
    s_base[w][lane] = inc - c;  // first rank of the bank in the row's bank-sorted order
    __syncwarp();
    for (int t0 = 0; t0 < L; t0 += 32) {
      const int t = t0 + lane;
      const bool valid = t < L;
      const int col = valid ? __ldg(Bj + b + t) : 0;
      const T val = valid ? __ldg(Bx + b + t) : T(0);
      const int bank = col & (BANKS - 1);
      const unsigned peers = __match_any_sync(FULL, valid ? bank : 32 + lane);
      const int r_in = __popc(peers & lt);
      int sidx = 0;
      if (valid) sidx = s_base[w][bank] + r_in;
      __syncwarp();
      if (valid && r_in == __popc(peers) - 1) s_base[w][bank] = sidx + 1;
      __syncwarp();
      if (valid) {
        PBEnt<T> e;
        e.col = col;
        e.val = val;
        int q = __float2int_rz((float)sidx * rG);  // sidx / G without the integer division (sidx < 2^24)
        int g = sidx - q * G;
        if (g >= G) g -= G, ++q;
        if (g < 0) g += G, --q;
        row[g * 32 + q] = e;  // dealt: consecutive ranks go to consecutive groups
      }
    }
    // the slots no entry was dealt to: group g holds ceil((L - g) / G) entries = L / G, one more for g < L % G
    const int qL = L / G, rL = L - qL * G;
    for (int g = 0; g < G; ++g) {
      if (lane >= qL + (g < rL)) {
        PBEnt<T> e;
        e.col = -1;
        e.val = T(0);
        row[(g << 5) + lane] = e;
      }
    }
  }
}

// Staging of k_num_dense_cta, split in two arrays so that nothing but the loaded (column, value)
// pairs rides in registers across the unrolled steps: the start of a piece is needed when its loads
// are issued (D steps ahead), its length and A value when it is applied.  (One 16-byte entry read
// with LDS.128 kept length and A value of all D pieces in flight in registers: 16 register moves per
// 8 steps at the loop end - a tenth of the loop's instructions - and 2 wavefronts per read as well.)
template <typename T> struct __align__(8) LenA {
  int len;
  T a;
};

// The pieces staged in s_start / s_la [0 .. cnt) applied to the accumulator in order, one barrier
// per piece.  Software pipeline: the (column, value) loads of pieces q+1 .. q+D are in flight while
// piece q is applied (the L2 round trip is ~5 barrier steps long), and the staging entries of the
// next fetch and the next apply are read one step early, so nothing but the accumulator read itself
// waits on shared memory.  The thread's index into a piece is clamped to the last entry of B, not to
// the piece: lanes past its end load a neighbour's entry and discard it.
template <typename T, int D>
__device__ __forceinline__ void dense_apply_staged(T *__restrict__ acc, const int *s_start, const LenA<T> *s_la, int cnt,
                                                   int tid, const PBEnt<T> *__restrict__ pb, int last) {
  int j[D];
  T b[D];
  int nstart = s_start[0];
#define DENSE_FETCH(S, Q)                      \
  {                                            \
    const int idx_ = min(nstart + tid, last);  \
    nstart = s_start[(Q) + 1];                 \
    ld_pb(pb + idx_, j[S], b[S]);              \
  }
#pragma unroll
  for (int s = 0; s < D; ++s) DENSE_FETCH(s, s)
  LenA<T> cu = s_la[0];
  for (int q = 0; q < cnt; q += D) {
#pragma unroll
    for (int s = 0; s < D; ++s) {
      // the accumulator read goes first; the next fetch is issued underneath its latency
      const int len = cu.len;
      const T a = cu.a;
      cu = s_la[q + s + 1];
      const bool piece = len != 0;  // CTA-uniform; only the padding after the last piece is empty
      const bool act = tid < len && j[s] >= 0;  // inside the piece and not one of the row's padding slots
      T *slot = acc + j[s];
      T cur = T(0);
      if (act) cur = *slot;
      const T prod = add_rn(mul_rn(a, b[s]), T(0));  // -0.0 -> +0.0
      DENSE_FETCH(s, q + D + s)
      if (act) *slot = add_rn(cur, prod);
      if (piece) __syncthreads();
    }
  }
#undef DENSE_FETCH
}

// Four consecutive accumulator columns per lane (one or two LDS.128).
template <typename T> struct Vec4;
template <> struct Vec4<float> {
  static __device__ __forceinline__ void load(const float *p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4 *>(p);
    v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
  }
  static __device__ __forceinline__ void fill(float *p, float x) { *reinterpret_cast<float4 *>(p) = make_float4(x, x, x, x); }
};
template <> struct Vec4<double> {
  static __device__ __forceinline__ void load(const double *p, double (&v)[4]) {
    const double2 t = *reinterpret_cast<const double2 *>(p), u = *reinterpret_cast<const double2 *>(p + 2);
    v[0] = t.x, v[1] = t.y, v[2] = u.x, v[3] = u.y;
  }
  static __device__ __forceinline__ void fill(double *p, double x) {
    *reinterpret_cast<double2 *>(p) = make_double2(x, x);
    *reinterpret_cast<double2 *>(p + 2) = make_double2(x, x);
  }
};

// Touched columns of one warp's range [c_lo, c_hi) (multiples of 128; the accumulator is padded
// with untouched columns).
template <typename T>
__device__ __forceinline__ int dense_count(const T *acc, int c_lo, int c_hi, int lane) {
  int mine = 0;
  for (int c = c_lo + 4 * lane; c < c_hi; c += 128) {
    T v[4];
    Vec4<T>::load(acc + c, v);
#pragma unroll
    for (int e = 0; e < 4; ++e) mine += !NegZero<T>::is(v[e]);
  }
  return warp_sum(mine);
}

// Ordered emit of that range, 32 columns per step (fully coalesced stores: the better choice for
// well-filled rows).  `steps` = the trip count, a CTA-uniform upper bound of (c_hi - c_lo) / 32 that
// the compiler can see is uniform (kernel parameters and blockDim only): with a loop bound derived
// from the warp index it cannot prove the warp converged at the vote and wraps every ballot in a
// BRA.DIV / BSSY / BSYNC sequence; and the stores are predicated on values computed outside the
// `if`, so no branch is left (24.5 -> 14 instructions per 32 columns; the sweep was a third of all
// instructions of the kernel on C3).
// The three stores of one emitted entry under ONE predicate, without a branch (ptxas turns the `if`
// into BSSY / BRA / BSYNC and sinks the address arithmetic into it).
__device__ __forceinline__ void emit_if(bool p, int *cj, int col, float *cx, float v, float *slot, float unt) {
  asm volatile(
      "{\n.reg .pred q;\nsetp.ne.s32 q, %0, 0;\n@q st.global.s32 [%1], %2;\n@q st.global.f32 [%3], %4;\n@q st.shared.f32 [%5], %6;\n}\n" ::"r"((int)p),
      "l"(cj), "r"(col), "l"(cx), "f"(v), "r"((unsigned)__cvta_generic_to_shared(slot)), "f"(unt)
      : "memory");
}
__device__ __forceinline__ void emit_if(bool p, int *cj, int col, double *cx, double v, double *slot, double unt) {
  asm volatile(
      "{\n.reg .pred q;\nsetp.ne.s32 q, %0, 0;\n@q st.global.s32 [%1], %2;\n@q st.global.f64 [%3], %4;\n@q st.shared.f64 [%5], %6;\n}\n" ::"r"((int)p),
      "l"(cj), "r"(col), "l"(cx), "d"(v), "r"((unsigned)__cvta_generic_to_shared(slot)), "d"(unt)
      : "memory");
}

template <typename T, bool SCALE>
__device__ __forceinline__ void dense_emit_ballot(T *acc, int c_lo, int c_hi, int steps, int lane, int *__restrict__ Cj,
                                                  T *__restrict__ Cx, int pos, T alpha, int col0 = 0) {
  const T UNT = NegZero<T>::value();
  const unsigned below = (1u << lane) - 1u;
  int c = c_lo + lane;
#pragma unroll 4
  for (int it = 0; it < steps; ++it, c += 32) {
    const bool in = c < c_hi;
    const T v = in ? acc[c] : UNT;
    const bool touched = !NegZero<T>::is(v);
    const unsigned bal = __ballot_sync(FULL, touched);
    const int o = pos + __popc(bal & below);
    const T out = SCALE ? mul_rn(alpha, v) : v;
    emit_if(touched, Cj + o, col0 + c, Cx + o, out, acc + c, UNT);
    pos += __popc(bal);
  }
}

// Ordered emit of that range, 128 columns per step: lane l owns columns c+4l .. c+4l+3, a warp
// scan of the per-lane counts gives its output offset; the accumulator is reset on the way.
template <typename T, bool SCALE>
__device__ __forceinline__ void dense_emit(T *acc, int c_lo, int c_hi, int steps, int lane, int *__restrict__ Cj,
                                           T *__restrict__ Cx, int pos, T alpha, int col0 = 0) {
  const T UNT = NegZero<T>::value();
  int c = c_lo + 4 * lane;
  for (int it = 0; it < steps; ++it, c += 128) {  // uniform trip count (see dense_emit_ballot): the shuffles stay on the converged path
    T v[4] = {UNT, UNT, UNT, UNT};
    if (c < c_hi) Vec4<T>::load(acc + c, v);
    bool t[4];
    int cnt = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      t[e] = !NegZero<T>::is(v[e]);
      cnt += t[e];
    }
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(FULL, inc, o);
      if (lane >= o) inc += u;
    }
    int o = pos + inc - cnt;
    if (cnt) {
      Vec4<T>::fill(acc + c, UNT);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (t[e]) {
          Cj[o] = col0 + c + e;
          Cx[o] = SCALE ? mul_rn(alpha, v[e]) : v[e];
          ++o;
        }
      }
    }
    pos += __shfl_sync(FULL, inc, 31);
  }
}

// ---- single-pass mode (SINGLE = true) ---------------------------------------------------------
// The same kernel without a structure pass in front of it: rows are taken in row order from a
// ticket counter, a row's output position is the number of entries of all earlier rows, obtained
// with a decoupled look-back over per-row status words (the chained-scan protocol of single-pass
// prefix sums: a row publishes its own count as soon as it knows it, then sums its predecessors'
// counts back to the nearest published inclusive prefix), and the kernel writes C.indptr itself.
// C.indices / C.data have room for an upper bound of nnz(C); the caller trims them afterwards.
// Every CTA of the grid must be resident (the host sizes the grid from the occupancy query): a row
// only ever waits for rows with smaller tickets, which are running or done.
struct SinglePassArgs {
  int *ticket;                  // zeroed before the launch
  unsigned long long *status;   // m words, zeroed before the launch: flag << 62 | value
  int *Cp;                      // m + 1 row pointers, written here
  unsigned long long *nnz_out;  // total, written by the last row
  const int *guard;             // optional: two flags (row-product overflow, B not canonical) set by the estimate
                                // kernels in front of this launch; the kernel does nothing when either is set
};
#define SP_FLAG_AGG (1ull << 62)
#define SP_FLAG_PREFIX (2ull << 62)
#define SP_VALUE_MASK ((1ull << 62) - 1)

// Warp-wide: publishes `own` for `row` and returns the number of entries of rows [0, row).
__device__ __forceinline__ long long single_pass_lookback(volatile unsigned long long *status, int row, long long own, int lane) {
  if (row == 0) {
    if (lane == 0) status[0] = SP_FLAG_PREFIX | (unsigned long long)own;
    return 0;
  }
  if (lane == 0) status[row] = SP_FLAG_AGG | (unsigned long long)own;
  long long excl = 0;
  for (int j = row - 1;; j -= 32) {  // lane l looks at row j - l (rows before 0 count as prefix 0)
    const int idx = j - lane;
    unsigned long long s;
    int fp;
    for (;;) {
      s = idx >= 0 ? status[idx] : SP_FLAG_PREFIX;
      const unsigned flag = (unsigned)(s >> 62);
      const unsigned pref = __ballot_sync(FULL, flag == 2u), inval = __ballot_sync(FULL, flag == 0u);
      fp = pref ? __ffs(pref) - 1 : 32;  // nearest predecessor with an inclusive prefix
      const unsigned upto = fp < 32 ? (2u << fp) - 1u : FULL;
      if ((inval & upto) == 0) break;
      __nanosleep(40);
    }
    long long v = lane <= fp ? (long long)(s & SP_VALUE_MASK) : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    excl += v;
    if (fp < 32) break;
  }
  if (lane == 0) status[row] = SP_FLAG_PREFIX | (unsigned long long)(excl + own);
  return excl;
}

template <typename T, bool SINGLE, int DEPTH>
__global__ void __launch_bounds__(DENSE_MAX_NT, 3)
k_num_dense_cta(const int *__restrict__ rows, int nrows, int n, const int *__restrict__ Ap,
                const int *__restrict__ Aj, const T *__restrict__ Ax, const int *__restrict__ Bp,
                const PBEnt<T> *__restrict__ pb, int pb_last, const int *__restrict__ Cp,
                int *__restrict__ Cj, T *__restrict__ Cx, T alpha, SinglePassArgs sp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];  // round_up(n, 128) accumulator columns
  __shared__ int s_start[DENSE_CAP + DENSE_PAD];             // the staged pieces of the round: first entry in B ...
  __shared__ LenA<T> s_la[DENSE_CAP + DENSE_PAD];            // ... length and A value
  __shared__ int s_wcnt[DENSE_MAX_NT / 32];
  __shared__ int s_scan[DENSE_MAX_NT / 32];
  __shared__ int s_row;
  __shared__ long long s_pos;
  T *acc = reinterpret_cast<T *>(smem_raw);
  const int NT = blockDim.x, tid = threadIdx.x, lane = lane_id(), w = tid >> 5, nw = NT >> 5;
  const T UNT = NegZero<T>::value();
  const int npad = (n + 127) & ~127;
  if (SINGLE && sp.guard && (sp.guard[0] | sp.guard[1])) return;  // grid-uniform
  for (int c = tid; c < npad; c += NT) acc[c] = UNT;
  // columns per warp for the output sweep, a multiple of 128
  const int cpw = ((npad + nw - 1) / nw + 127) & ~127;
  const int c_lo = min(w * cpw, npad), c_hi = min(c_lo + cpw, npad);
  __syncthreads();

  for (int ridx = blockIdx.x;; ridx += gridDim.x) {
    int row;
    if (SINGLE) {
      // The ticket is drawn when the CTA is free, not ahead of time: a row waits for every row before it, so
      // tickets must go out in the order in which CTAs come free.  Measured on C3 (kernel time): drawn here
      // 1.22 ms; drawn a row ahead 1.98 ms (a late CTA sits on a low ticket and all rows behind it spin);
      // drawn before the count pass 1.36 ms.
      if (tid == 0) s_row = atomicAdd(sp.ticket, 1);
      __syncthreads();  // the previous read of s_row lies before the barriers of the row body
      row = s_row;
      if (row >= nrows) break;
    } else {
      if (ridx >= nrows) break;
      row = rows[ridx];
    }
    const int a_begin = __ldg(Ap + row), a_end = __ldg(Ap + row + 1);
    for (int base = a_begin; base < a_end; base += NT) {
      // ---- stage: thread t takes A nonzero base+t; its B row is cut into pieces of <= NT
      //      entries; an exclusive scan of the piece counts gives every piece its slot, in A order
      const int t = base + tid;
      int start = 0, len = 0;
      T av = T(0);
      if (t < a_end) {
        const int kk = __ldg(Aj + t);
        const int bp = __ldg(Bp + kk);
        start = pb_start(bp, kk);
        len = (__ldg(Bp + kk + 1) - bp + 31) & ~31;  // the row in the packed copy: whole groups of 32 entries
        av = __ldg(Ax + t);
      }
      const int pieces = (len + NT - 1) / NT;  // 0 for an empty B row: it gets no slot at all
      int inc = pieces;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += u;
      }
      if (lane == 31) s_scan[w] = inc;
      __syncthreads();
      int off = inc - pieces, total = 0;
      for (int q = 0; q < nw; ++q) {
        const int v = s_scan[q];
        if (q < w) off += v;
        total += v;
      }
      for (int r0 = 0; r0 < total; r0 += DENSE_CAP) {
        const int cnt = min(DENSE_CAP, total - r0);
        const int p_lo = max(0, r0 - off), p_hi = min(pieces, r0 + DENSE_CAP - off);
        for (int pc = p_lo; pc < p_hi; ++pc) {
          LenA<T> la;
          la.len = min(NT, len - pc * NT);
          la.a = av;
          s_start[off + pc - r0] = start + pc * NT;
          s_la[off + pc - r0] = la;
        }
        if (tid < DENSE_PAD) {
          LenA<T> z;
          z.len = 0;
          z.a = T(0);
          s_start[cnt + tid] = 0;
          s_la[cnt + tid] = z;
        }
        __syncthreads();
        dense_apply_staged<T, DEPTH>(acc, s_start, s_la, cnt, tid, pb, pb_last);
        __syncthreads();
      }
      if (total == 0) __syncthreads();  // s_scan is rewritten by the next chunk
    }
    // ---- output: warp w owns columns [c_lo, c_hi); count, prefix over the warps, ordered emit ----
    const int mine = a_begin < a_end ? dense_count<T>(acc, c_lo, c_hi, lane) : 0;
    if (lane == 0) s_wcnt[w] = mine;
    __syncthreads();
    int pos, row_nnz;
    if (SINGLE) {
      row_nnz = 0;
      for (int q = 0; q < nw; ++q) row_nnz += s_wcnt[q];
      if (w == 0) {
        const long long excl = single_pass_lookback(sp.status, row, row_nnz, lane);
        if (lane == 0) {
          s_pos = excl;
          sp.Cp[row] = (int)excl;
          if (row == nrows - 1) {
            sp.Cp[nrows] = (int)(excl + row_nnz);
            *sp.nnz_out = (unsigned long long)(excl + row_nnz);
          }
        }
      }
      __syncthreads();
      pos = (int)s_pos;
      if (row_nnz == 0) continue;  // CTA-uniform; the next ticket's barrier orders the reuse of s_pos / s_wcnt
    } else {
      pos = __ldg(Cp + row);
      row_nnz = __ldg(Cp + row + 1) - pos;
    }
    for (int q = 0; q < w; ++q) pos += s_wcnt[q];
    // well-filled rows: 32 columns per step with fully coalesced stores; sparse rows: 128 columns
    // per step (a quarter of the instructions, stores strided by the fill)
    if (row_nnz > (npad >> 2)) {
      if (alpha == T(1)) dense_emit_ballot<T, false>(acc, c_lo, c_hi, cpw >> 5, lane, Cj, Cx, pos, alpha);
      else dense_emit_ballot<T, true>(acc, c_lo, c_hi, cpw >> 5, lane, Cj, Cx, pos, alpha);
    } else {
      if (alpha == T(1)) dense_emit<T, false>(acc, c_lo, c_hi, cpw >> 7, lane, Cj, Cx, pos, alpha);
      else dense_emit<T, true>(acc, c_lo, c_hi, cpw >> 7, lane, Cj, Cx, pos, alpha);
    }
    __syncthreads();
  }
}

// ---- wide matrices (n * sizeof(T) > shared memory): column blocks -----------------------------
// The n columns are cut into nblk blocks of W accumulators; for each block the CTA walks the A
// row once and applies only the part of every B row that falls into the block (B rows are
// sorted, so that part is one segment), then sweeps the block.  Blocks are processed left to
// right, so the row's output is appended block after block: sorted, no hash table, no sort, no
// global-memory accumulator.  Requires canonical B (the host checks).
// bounds[kk*(nblk+1) + b] = first entry of B row kk with column >= b*W (b = 0..nblk; the last is
// the row end): the segment of row kk inside column block b is [bounds[..b], bounds[..b+1]).
// Built once per plan and value type; k*(nblk+1) ints of workspace.
__global__ void __launch_bounds__(256)
k_block_bounds(int k, int nblk, int W, const int *__restrict__ Bp, const int *__restrict__ Bj,
               int *__restrict__ bounds) {
  const long long total = (long long)k * (nblk + 1);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int kk = (int)(i / (nblk + 1)), b = (int)(i % (nblk + 1));
    int lo = __ldg(Bp + kk), hi = __ldg(Bp + kk + 1);
    if (b == nblk) lo = hi;
    else if (b > 0) {
      const int key = b * W;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(Bj + mid) < key) lo = mid + 1;
        else hi = mid;
      }
    }
    bounds[i] = lo;
  }
}

// k_num_dense_pieces: one B-row PIECE per barrier, pipelined.
// The heavy rows of power-law inputs multiply A entries that point at hub rows of B, so their
// segments inside a column block are long (R-MAT scale 16 / 18: 130 / 105 products per piece of
// at most 256): the scheme of k_num_dense_cta - one piece of one B row per barrier step, columns
// inside a piece distinct, order = A-row order for free, (column, value) loads 16 pieces ahead -
// needs no search, no conflict handling and no retry.  Two earlier designs that flattened the
// segments into a product stream (a claim table with CTA-wide retries; owner warps fed by a
// stable counting sort) spent 10 warp instructions per product on search and ordering and were
// 1.4-1.9x slower at scale 16-18 (profiles/r01_heavy_row_kernels.md).
// Touched columns are recorded in a bitmap (one ATOMS.OR on the first touch of a column, found by
// the -0.0 marker), so the per-block sweep reads W/32 words instead of W accumulators: lane l
// takes word l of a group of 32, a warp scan of the popcounts gives its output offset, and it
// walks its own set bits (sorted output; sparse rows pay for their entries, not for the block
// width).
constexpr int BLOCKED_DEPTH = 16;             // pieces whose loads are in flight (an L2 round trip is ~8 barrier steps)
constexpr int BLOCKED_PAD = 2 * BLOCKED_DEPTH;  // zero-length staging entries after the last piece

template <typename T, int D>
__device__ __forceinline__ void blocked_apply_staged(T *__restrict__ acc, unsigned *__restrict__ bm,
                                                     const DenseSeg<T> *s_seg, int cnt, int tid, int c0,
                                                     const int *__restrict__ Bj, const T *__restrict__ Bx, int last) {
  // only the loaded (column, value) pairs ride in registers; length and A value of a piece are
  // read back from the staging entry one step before its turn
  int j[D];
  T b[D];
  DenseSeg<T> nx = s_seg[0];
#define BLOCKED_FETCH(S, Q)                     \
  {                                             \
    const int idx_ = min(nx.start + tid, last); \
    nx = s_seg[(Q) + 1];                        \
    j[S] = ld_inplace(Bj + idx_);               \
    b[S] = ld_inplace(Bx + idx_);               \
  }
#pragma unroll
  for (int s = 0; s < D; ++s) BLOCKED_FETCH(s, s)
  DenseSeg<T> cu = s_seg[0];
  for (int q = 0; q < cnt; q += D) {
#pragma unroll
    for (int s = 0; s < D; ++s) {
      const int len = cu.len;
      const T a = cu.a;
      cu = s_seg[q + s + 1];
      const bool piece = len != 0;  // CTA-uniform; only the padding after the last piece is empty
      const bool act = tid < len;
      const int lc = j[s] - c0;
      T cur = T(0);
      if (act) cur = acc[lc];
      const T prod = add_rn(mul_rn(a, b[s]), T(0));  // -0.0 -> +0.0
      BLOCKED_FETCH(s, q + D + s)
      if (act) {
        acc[lc] = add_rn(cur, prod);
        if (NegZero<T>::is(cur)) atomicOr(bm + (lc >> 5), 1u << (lc & 31));
      }
      if (piece) __syncthreads();
    }
  }
#undef BLOCKED_FETCH
}

// mbarrier hand-off between the warps of a CTA (one arrival per phase): the release/acquire pair
// orders the accumulator updates of consecutive windows of the product stream.
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, int parity) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(bar);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(a), "r"(parity) : "memory");
}

// acc[col] += v for the active lanes of one window of the product stream; lanes that hold the same
// column (products of different B rows) are chained in lane order (= stream order) by the lowest
// lane.  `peers` = __match_any_sync over the columns (computed before the warp's turn).
template <typename T>
__device__ __forceinline__ void window_apply(T *__restrict__ acc, unsigned *__restrict__ bm, int col, bool act, T v,
                                             unsigned peers, bool any_multi, int maxn) {
  const int lane = lane_id();
  const bool lead = act && (peers & ((1u << lane) - 1u)) == 0;
  T cur = T(0);
  if (lead) cur = acc[col];
  T a = add_rn(cur, v);
  if (any_multi) {  // warp-uniform
    unsigned m = peers & (peers - 1);  // the members after the leader, lowest lane first
    for (int r = 1; r < maxn; ++r) {
      const int src = m ? __ffs(m) - 1 : lane;
      const T vv = __shfl_sync(FULL, v, src);
      if (lead && m) a = add_rn(a, vv);
      m &= m - 1;
    }
  }
  if (lead) {
    acc[col] = a;
    if (NegZero<T>::is(cur)) atomicOr(bm + (col >> 5), 1u << (col & 31));
  }
}

// Chunks with fewer products per piece than this take the window path: a window of 32 products
// costs ~420 cycles of hand-off chain, a piece ~340 cycles of barrier step (measured: R-MAT scale
// 16/18 unchanged, scale 20 961 -> 827 ms, scale 22 16.3 -> 12.0 s; 64 and above lose at scale 16-18).
// ---- hub rows split over CTAs ------------------------------------------------------------------
// The heaviest rows of a power-law matrix are a whole CTA's work for tens of milliseconds (R-MAT scale
// 20: the top row holds 9.7e6 products, ~60 ms of one CTA), which is what a strong-scaled run on 8 GPUs
// has left per rank: the longest row, not the work, then sets the time.  The rows within a factor 4 of
// the largest (the two top buckets of the work-ordered list, at most HUB_MAX rows) are therefore cut
// into their column blocks - one task per (row, block), taken by different CTAs.  A block's entries
// start after those of the row's earlier blocks, so their counts are needed first: k_hub_counts
// walks the structure of every (hub row, block) once (bitmap in shared memory, ~5 % of the products),
// k_hub_offsets turns the counts of a row into offsets.
constexpr int HUB_MAX = 256;

__device__ __forceinline__ int hub_rows(const int *__restrict__ hist, int nrows) {
  return min(min(hist[15] + hist[14], HUB_MAX), nrows);
}

__global__ void __launch_bounds__(256)
k_hub_counts(const int *__restrict__ rows, int nrows, const int *__restrict__ hist, int *__restrict__ ticket, int W, int nblk,
             const int *__restrict__ Ap, const int *__restrict__ Aj, const int *__restrict__ bounds,
             const int *__restrict__ Bj, int *__restrict__ counts /* [HUB_MAX][nblk + 1] */) {
  extern __shared__ unsigned hub_bm[];  // W / 32 words
  __shared__ int s_t, s_cnt;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, nw = blockDim.x >> 5;
  const int H = hub_rows(hist, nrows), ntasks = H * nblk, nwords = W >> 5;
  for (;;) {
    if (tid == 0) {
      s_t = atomicAdd(ticket, 1);
      s_cnt = 0;
    }
    for (int i = tid; i < nwords; i += blockDim.x) hub_bm[i] = 0u;
    __syncthreads();
    const int t = s_t;
    if (t >= ntasks) break;
    const int hrow = t / nblk, b = t - hrow * nblk, row = rows[hrow], c0 = b * W;
    const int a_end = __ldg(Ap + row + 1);
    for (int e = __ldg(Ap + row) + w; e < a_end; e += nw) {
      const int *bd = bounds + (size_t)__ldg(Aj + e) * (nblk + 1) + b;
      const int se = __ldg(bd + 1);
      for (int u = __ldg(bd) + lane; u < se; u += 32) {
        const int lc = __ldg(Bj + u) - c0;
        atomicOr(hub_bm + (lc >> 5), 1u << (lc & 31));
      }
    }
    __syncthreads();
    int c = 0;
    for (int i = tid; i < nwords; i += blockDim.x) c += __popc(hub_bm[i]);
    c = warp_sum(c);
    if (lane == 0 && c) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (tid == 0) counts[hrow * (nblk + 1) + b] = s_cnt;
    __syncthreads();
  }
}

// one warp per hub row: counts[h][0 .. nblk) -> exclusive prefix in place (counts[h][nblk] = the row's total)
__global__ void __launch_bounds__(256)
k_hub_offsets(const int *__restrict__ hist, int nrows, int nblk, int *__restrict__ counts) {
  const int lane = threadIdx.x & 31;
  const int h = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (h >= hub_rows(hist, nrows)) return;
  int *c = counts + h * (nblk + 1);
  int run = 0;
  for (int b0 = 0; b0 < nblk; b0 += 32) {
    const int b = b0 + lane;
    const int v = b < nblk ? c[b] : 0;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(FULL, inc, o);
      if (lane >= o) inc += u;
    }
    if (b < nblk) c[b] = run + inc - v;
    run += __shfl_sync(FULL, inc, 31);
  }
  if (lane == 0) c[nblk] = run;
}

constexpr int PIECES_FLAT_BELOW = 32;

template <typename T, int NT, int WPT>
__global__ void __launch_bounds__(NT, 1)
k_num_dense_pieces(const int *__restrict__ rows, int nrows, int *__restrict__ ticket, int n, int W, int nblk,
                   const int *__restrict__ Ap, const int *__restrict__ Aj, const T *__restrict__ Ax,
                   const int *__restrict__ bounds, const int *__restrict__ Bj, const T *__restrict__ Bx,
                   int nnzB, const int *__restrict__ Cp, int *__restrict__ Cj, T *__restrict__ Cx, T alpha, int flat_below,
                   const int *__restrict__ hub_hist, const int *__restrict__ hub_offs) {
  constexpr int NW = NT / 32, CAP = 2 * NT;
  extern __shared__ __align__(16) unsigned char smem_raw[];  // W accumulators, then W/32 bitmap words
  __shared__ DenseSeg<T> s_seg[CAP + BLOCKED_PAD];
  __shared__ int s_scan[NW];
  __shared__ int s_scanp[NW];  // products per warp of the staged chunk
  __shared__ int s_gcnt[64];   // touched columns per group of 1024 (W <= 65536)
  __shared__ int s_ticket;
  __shared__ __align__(8) unsigned long long s_turn[NW];  // hand-off barriers of the window path
  // the window path reuses the staging array: prefix of the segment lengths, start - prefix, A value
  static_assert(sizeof(s_seg) >= (2 * NT + 2) * sizeof(int) + 8 + NT * sizeof(T), "staging alias");
  static_assert(sizeof(s_seg) >= NW * 64 * sizeof(int), "sweep alias");
  int *f_pre = reinterpret_cast<int *>(s_seg);
  int *f_sbase = f_pre + NT + 1;
  T *f_a = reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(s_seg) + (((2 * NT + 1) * sizeof(int) + 7) & ~size_t(7)));
  T *acc = reinterpret_cast<T *>(smem_raw);
  unsigned *bm = reinterpret_cast<unsigned *>(smem_raw + (size_t)W * sizeof(T));
  const int tid = threadIdx.x, lane = lane_id(), w = tid >> 5;
  if (tid < NW) mbar_init(&s_turn[tid], 1);
  int gwin = 0;   // turns (groups of WPT windows) handed out so far (CTA-uniform): turn g belongs to warp g % NW
  int turns = 0;  // hand-offs this warp has consumed (phase of its barrier)
  const int nwords = W >> 5;  // W is a multiple of 128
  const T UNT = NegZero<T>::value();
  for (int c = tid; c < W; c += NT) acc[c] = UNT;
  for (int c = tid; c < nwords; c += NT) bm[c] = 0u;
  const int ngrp = (nwords + 31) >> 5;
  __syncthreads();
  // rows are handed out by a global ticket: heavy rows differ 1000-fold in work, a fixed
  // round-robin deal left the slowest CTA 1.2-1.35x behind the mean (R-MAT scale 16-18)
  // hub_hist != nullptr: the first H rows of the (work-ordered) list are hub rows, cut into H * nblk tasks of one
  // column block each, whose offsets inside the row are in hub_offs (k_hub_counts / k_hub_offsets); tickets
  // [0, H * nblk) are those tasks (row-major, heaviest rows first), the later ones the remaining rows
  const int H = hub_hist ? hub_rows(hub_hist, nrows) : 0;
  const int hub_tasks = H * nblk;
  while (true) {
    if (tid == 0) s_ticket = atomicAdd(ticket, 1);
    __syncthreads();
    const int tk = s_ticket;  // the next write is behind the barriers of this row's blocks
    int ridx, cb_lo = 0, cb_hi = nblk, hub_off = 0;
    if (tk < hub_tasks) {
      ridx = tk / nblk;
      cb_lo = tk - ridx * nblk;
      cb_hi = cb_lo + 1;
      hub_off = __ldg(hub_offs + ridx * (nblk + 1) + cb_lo);
    } else {
      ridx = tk - hub_tasks + H;
    }
    if (ridx >= nrows) break;
    const int row = rows[ridx];
    const int a_begin = __ldg(Ap + row), a_end = __ldg(Ap + row + 1);
    int out = __ldg(Cp + row) + hub_off;
    for (int cb = cb_lo; cb < cb_hi; ++cb) {
      const int c0 = cb * W;
      for (int base = a_begin; base < a_end; base += NT) {
        // ---- stage: thread t takes A nonzero base+t; the segment of its B row inside the block is
        //      cut into pieces of <= NT entries, slots from an exclusive scan of the piece counts
        const int t = base + tid;
        int start = 0, len = 0;
        T av = T(0);
        if (t < a_end) {
          const int kk = __ldg(Aj + t);
          const int *bd = bounds + (size_t)kk * (nblk + 1) + cb;
          start = __ldg(bd);
          len = __ldg(bd + 1) - start;
          av = __ldg(Ax + t);
        }
        const int pieces = (len + NT - 1) / NT;
        int inc = pieces;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int u = __shfl_up_sync(FULL, inc, o);
          if (lane >= o) inc += u;
        }
        int incp = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int u = __shfl_up_sync(FULL, incp, o);
          if (lane >= o) incp += u;
        }
        if (lane == 31) {
          s_scan[w] = inc;
          s_scanp[w] = incp;
        }
        __syncthreads();
        int off = inc - pieces, total = 0, poff = incp - len, P = 0;
#pragma unroll
        for (int q = 0; q < NW; ++q) {
          const int v = s_scan[q], vp = s_scanp[q];
          if (q < w) off += v, poff += vp;
          total += v, P += vp;
        }
        if (total > 0 && P < flat_below * total) {
          // ---- window path: short segments.  The chunk's products form one stream (A order);
          //      window i = products [32 i, 32 i + 32) is applied by ONE warp, same-column lanes
          //      chained in lane order; the windows are applied strictly one after the other
          //      (warp g % NW takes window g and is handed the turn by its predecessor through an
          //      mbarrier), while location, loads and the column match of the following windows
          //      run ahead on the other warps.
          f_pre[tid] = poff;
          f_sbase[tid] = start - poff;
          f_a[tid] = av;
          if (tid == 0) f_pre[NT] = P;
          __syncthreads();
          // a TURN = WPT consecutive windows applied by one warp: the hand-off (~420 cycles) is paid
          // once per turn, the windows inside a turn follow each other in program order
          const int nwin = (P + 31) >> 5;
          const int nturn = (nwin + WPT - 1) / WPT;
          int i = (w - gwin) & (NW - 1);  // my first turn of this chunk
          int col[WPT];
          T prod[WPT];
          unsigned actm = 0;  // bit u: my lane holds a product in window u of the turn
          auto fetch = [&](int turn) {
            actm = 0;
#pragma unroll
            for (int u = 0; u < WPT; ++u) {
              const int pp = ((turn * WPT + u) << 5) + lane;
              col[u] = 0;
              prod[u] = T(0);
              if (turn < nturn && pp < P) {
                actm |= 1u << u;
                int lo = 0, hi = NT;  // largest q with f_pre[q] <= pp
#pragma unroll
                for (int st = 0; st < 8; ++st) {
                  const int mid = (lo + hi) >> 1;
                  if (f_pre[mid] <= pp) lo = mid;
                  else hi = mid;
                }
                const int idx = f_sbase[lo] + pp;
                col[u] = __ldg(Bj + idx) - c0;
                prod[u] = add_rn(mul_rn(f_a[lo], __ldg(Bx + idx)), T(0));  // -0.0 -> +0.0
              }
            }
          };
          fetch(i);
          for (; i < nturn; i += NW) {
            int mycol[WPT];
            T myprod[WPT];
            const unsigned myact = actm;
#pragma unroll
            for (int u = 0; u < WPT; ++u) mycol[u] = col[u], myprod[u] = prod[u];
            fetch(i + NW);  // the next turn of this warp is located and requested before the wait
            unsigned peers[WPT];
            int maxn[WPT];
#pragma unroll
            for (int u = 0; u < WPT; ++u) {
              const bool a = (myact >> u) & 1u;
              peers[u] = __match_any_sync(FULL, a ? mycol[u] : (-1 - lane));
              const bool any_multi = __any_sync(FULL, a && (peers[u] & (peers[u] - 1)));
              maxn[u] = any_multi ? __reduce_max_sync(FULL, __popc(peers[u])) : 1;
            }
            if (gwin + i > 0) {  // every turn but the very first of the CTA waits for its predecessor
              mbar_wait(&s_turn[w], turns & 1);
              ++turns;
            }
#pragma unroll
            for (int u = 0; u < WPT; ++u) {
              if ((u << 5) < P - ((i * WPT) << 5)) {  // warp-uniform: window u of this turn exists
                window_apply<T>(acc, bm, mycol[u], (myact >> u) & 1u, myprod[u], peers[u], maxn[u] > 1, maxn[u]);
                __syncwarp();
              }
            }
            if (lane == 0) mbar_arrive(&s_turn[(w + 1) & (NW - 1)]);
          }
          gwin += nturn;
          __syncthreads();
          continue;
        }
        for (int r0 = 0; r0 < total; r0 += CAP) {
          const int cnt = min(CAP, total - r0);
          const int p_lo = max(0, r0 - off), p_hi = min(pieces, r0 + CAP - off);
          for (int pc = p_lo; pc < p_hi; ++pc) {
            DenseSeg<T> sg;
            sg.start = start + pc * NT;
            sg.len = min(NT, len - pc * NT);
            sg.a = av;
            s_seg[off + pc - r0] = sg;
          }
          if (tid < BLOCKED_PAD) {
            DenseSeg<T> z;
            z.start = 0;
            z.len = 0;
            z.a = T(0);
            s_seg[cnt + tid] = z;
          }
          __syncthreads();
          blocked_apply_staged<T, BLOCKED_DEPTH>(acc, bm, s_seg, cnt, tid, c0, Bj, Bx, nnzB - 1);
          __syncthreads();
        }
        if (total == 0) __syncthreads();  // s_scan is rewritten by the next chunk
      }
      // ---- sweep of the block through the bitmap.  Groups of 32 words (1024 columns) are dealt
      //      to the warps round-robin (power-law columns crowd the low indices: contiguous ranges
      //      would leave one warp with most of the entries); count per group, prefix, ordered emit
      for (int g = w; g < ngrp; g += NW) {
        const int i = (g << 5) + lane;
        const int c = warp_sum(i < nwords ? __popc(bm[i]) : 0);
        if (lane == 0) s_gcnt[g] = c;
      }
      __syncthreads();
      const int g0 = lane < ngrp ? s_gcnt[lane] : 0, g1 = lane + 32 < ngrp ? s_gcnt[lane + 32] : 0;
      int inc0 = g0, inc1 = g1;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u0 = __shfl_up_sync(FULL, inc0, o), u1 = __shfl_up_sync(FULL, inc1, o);
        if (lane >= o) inc0 += u0, inc1 += u1;
      }
      const int tot0 = __shfl_sync(FULL, inc0, 31), tot = tot0 + __shfl_sync(FULL, inc1, 31);
      const int ex0 = inc0 - g0, ex1 = tot0 + inc1 - g1;  // exclusive prefix of groups lane and lane + 32
      // Emit.  Every warp takes a contiguous run of groups holding about tot / NW entries (boundaries from the
      // prefix: the hot low columns no longer unbalance contiguous ranges) and walks it with a running output
      // offset.  Non-empty bitmap words are first gathered, 32 at a time, into the warp's pending list, and only
      // then handed one per lane: in the sparse stretches of a power-law row most words of a group are empty and
      // a per-group scan + walk spent ~3 warp instructions per entry on them.
      {
        // groups are assigned by the rank of their first entry: run w = the groups whose exclusive prefix lies in
        // [tot w / NW, tot (w + 1) / NW); groups_below(t) = number of groups with prefix < t = first group at or past t
        auto groups_below = [&](int t) {
          const unsigned b0 = __ballot_sync(FULL, lane < ngrp && ex0 < t);
          const unsigned b1 = __ballot_sync(FULL, lane + 32 < ngrp && ex1 < t);
          return __popc(b0) + __popc(b1);
        };
        const int gbeg = groups_below((int)((long long)tot * w / NW));
        const int gend = groups_below((int)((long long)tot * (w + 1) / NW));  // trailing empty groups belong to no run
        int run = out + __shfl_sync(FULL, gbeg < 32 ? ex0 : ex1, gbeg & 31);
        int *pend = reinterpret_cast<int *>(s_seg) + 64 * w;  // 64 pending words per warp; the staging array is idle during the sweep
        int npend = 0;
        const unsigned below = (1u << lane) - 1u;
        auto flush = [&](int c) {  // the first c (<= 32) pending words, one per lane
          unsigned word = 0u;
          int i = 0;
          if (lane < c) {
            i = pend[lane];
            word = bm[i];
            bm[i] = 0u;
          }
          const int cnt = __popc(word);
          int inc = cnt;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += u;
          }
          int *cj = Cj + (run + inc - cnt);
          T *cx = Cx + (run + inc - cnt);
          T *ac = acc + (i << 5);
          const int colbase = c0 + (i << 5);
          while (word) {
            const int bit = __ffs(word) - 1;
            word &= word - 1;
            const T v = ac[bit];
            *cx++ = alpha == T(1) ? v : mul_rn(alpha, v);
            ac[bit] = UNT;
            *cj++ = colbase + bit;
          }
          run += __shfl_sync(FULL, inc, 31);
        };
        for (int g = gbeg; g < gend; ++g) {
          const int i = (g << 5) + lane;
          const unsigned word = i < nwords ? bm[i] : 0u;
          const unsigned nz = __ballot_sync(FULL, word != 0u);
          if (nz == 0u) continue;  // warp-uniform
          if (word) pend[npend + __popc(nz & below)] = i;
          npend += __popc(nz);
          __syncwarp();
          if (npend >= 32) {
            flush(32);
            __syncwarp();
            const int rest = npend - 32;
            const int moved = lane < rest ? pend[32 + lane] : 0;
            __syncwarp();
            if (lane < rest) pend[lane] = moved;
            npend = rest;
            __syncwarp();
          }
        }
        if (npend > 0) flush(npend);
      }
      out += tot;
      __syncthreads();
    }
  }
}

}  // namespace b200sp
