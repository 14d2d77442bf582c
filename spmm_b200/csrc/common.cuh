// common.cuh -- device helpers shared by the SpGEMM / SpMV kernels of libb200sp (sm_100a).
//
// Everything on this path is irregular integer + fp32/fp64 scalar work bound by L2/HBM
// gathers and shared-memory accumulator traffic; there is no GEMM-shaped tile, so no
// tcgen05/TMEM.  The rules that matter: coalesced streaming of index/value
// arrays, warp shuffles/ballots/match for reductions and conflict handling, shared-memory
// accumulators sized per row bin, grids sized in multiples of the 148 SMs.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200sp {

constexpr unsigned FULL = 0xffffffffu;
constexpr int EMPTY = -1;  // empty hash slot (column indices are >= 0)


__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// Separately rounded multiply and add: the contract with scipy's csr_matmat
// (`sums[k] += v * Bx[kk]` compiled without FMA contraction).  The intrinsics stop nvcc from
// fusing them into an FMA.
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

// Multiplicative hash to a power-of-two table (mask = S-1).
__device__ __forceinline__ unsigned hash_col(int j, unsigned mask) {
  return ((unsigned)j * 0x9E3779B1u >> 7) & mask;
}

// Insert column j into an open-addressing set in shared memory; returns the slot.
// `fresh` is set when this call created the entry.  Safe under concurrent inserts of equal
// or different keys (atomicCAS decides).
__device__ __forceinline__ int hash_insert(int *keys, unsigned mask, int j, bool &fresh) {
  unsigned h = hash_col(j, mask);
  fresh = false;
  while (true) {
    int cur = keys[h];
    if (cur == j) return (int)h;
    if (cur == EMPTY) {
      int old = atomicCAS(&keys[h], EMPTY, j);
      if (old == EMPTY) {
        fresh = true;
        return (int)h;
      }
      if (old == j) return (int)h;
    }
    h = (h + 1) & mask;
  }
}

__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ long long warp_sum(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
template <int G>
__device__ __forceinline__ int group_sum(int v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o, G);
  return v;
}

__device__ __forceinline__ int next_pow2(int v) {
  return v <= 1 ? 1 : 1 << (32 - __clz(v - 1));
}

// -------------------------------------------------------------------------------------
// Product stream.  The intermediate products of one row of A, in scipy's order (A-row
// storage order, then B-row storage order), dealt out to a group of G lanes.  All 32 lanes of
// the warp must call this together (each group may own a different row, or an empty range).
//
// Two modes per chunk of G A-nonzeros, chosen warp-uniformly:
//   * balanced: a load-balanced search maps stream position p -> (A nonzero, offset) with
//     log2(G) shuffles, so rows of B with 1-8 entries still keep every lane busy.  Lanes of one
//     step may hold the same output column (from different B rows) -> `may_conflict = true`.
//   * row-at-a-time (G == 32 only): B rows long enough that striding one row with the whole
//     warp wastes little; columns inside one B row are distinct -> `may_conflict = false`.
//
// f(active, bidx, aidx, may_conflict) is called convergently by the full warp each step;
// bidx indexes Bj/Bx, aidx indexes Aj/Ax.  Step order == stream order, lane order == stream
// order inside a step.
// -------------------------------------------------------------------------------------
// load(act, bidx, aidx) issues the global loads of one step and returns them by value (a small
// struct of registers); use(act, loaded, may_conflict) consumes a step.  The loads of step s+1 are
// issued before step s is consumed (one step of software prefetch), so the L2 latency of the
// B gathers overlaps the hash/accumulate work instead of adding to every step.
template <int G, class LoadF, class UseF>
__device__ __forceinline__ void stream_products(int a_begin, int a_end, const int *__restrict__ Aj,
                                                const int *__restrict__ Bp, LoadF &&load, UseF &&use) {
  using V = decltype(load(true, 0, 0));
  const int lane = lane_id();
  const int sub = lane & (G - 1);
  for (int base = a_begin; __any_sync(FULL, base < a_end); base += G) {
    const int t = base + sub;
    int start = 0, len = 0;
    if (t < a_end) {
      const int kk = __ldg(Aj + t);
      start = __ldg(Bp + kk);
      len = __ldg(Bp + kk + 1) - start;
    }
    int incl = len;
#pragma unroll
    for (int o = 1; o < G; o <<= 1) {
      const int v = __shfl_up_sync(FULL, incl, o, G);
      if (sub >= o) incl += v;
    }
    const int total = __shfl_sync(FULL, incl, G - 1, G);
    const int excl = incl - len;
    bool row_mode = false;
    if (G == 32) {
      const int cnt = min(32, a_end - base);
      row_mode = total >= 24 * cnt;  // average B row >= 24 entries (warp-uniform for G == 32)
    }
    V cur;
    bool cur_act = false, have = false;  // `have` is warp-uniform
    if (row_mode) {
      const int cnt = min(32, a_end - base);
      for (int q = 0; q < cnt; ++q) {
        const int qs = __shfl_sync(FULL, start, q);
        const int ql = __shfl_sync(FULL, len, q);
        for (int t0 = 0; t0 < ql; t0 += 32) {
          const int tt = t0 + lane;
          const bool act = tt < ql;
          V nxt = load(act, qs + tt, base + q);
          if (have) use(cur_act, cur, false);
          cur = nxt;
          cur_act = act;
          have = true;
        }
      }
      if (have) use(cur_act, cur, false);
    } else {
      for (int p0 = 0; __any_sync(FULL, p0 < total); p0 += G) {
        const int p = p0 + sub;
        const bool act = p < total;
        int q = 0;  // number of group lanes whose inclusive offset is <= p
#pragma unroll
        for (int s = G / 2; s >= 1; s >>= 1) {
          const int v = __shfl_sync(FULL, incl, q + s - 1, G);
          if (v <= p) q += s;
        }
        const int qs = __shfl_sync(FULL, start, q, G);
        const int qe = __shfl_sync(FULL, excl, q, G);
        V nxt = load(act, qs + (p - qe), base + q);
        if (have) use(cur_act, cur, true);
        cur = nxt;
        cur_act = act;
        have = true;
      }
      if (have) use(cur_act, cur, true);
    }
  }
}

// Ordered accumulate of one step of the stream into shared-memory accumulators.
// Lanes that hit the same slot in this step are combined by the lowest lane in lane order
// (== stream order), one rounded add at a time, so the result equals the sequential sum.
// `slot` indexes this group's accumulator array `vals`; `uid` must identify the accumulator
// uniquely across the whole warp (groups own disjoint tables: uid = group * S + slot).
// Ends with __syncwarp so the next step sees the stores.
template <typename T>
__device__ __forceinline__ void ordered_add(T *vals_base, int slot_id, int uid, bool act, T v,
                                            bool may_conflict) {
  const int lane = lane_id();
  if (!may_conflict) {
    if (act) vals_base[slot_id] = add_rn(vals_base[slot_id], v);
    __syncwarp();
    return;
  }
  const unsigned peers = __match_any_sync(FULL, act ? uid : (-1 - lane));
  const bool multi = act && (peers & (peers - 1));
  if (!__any_sync(FULL, multi)) {
    if (act) vals_base[slot_id] = add_rn(vals_base[slot_id], v);
  } else {
    const int leader = __ffs(peers) - 1;
    const int maxn = __reduce_max_sync(FULL, __popc(peers));
    T acc = act ? vals_base[slot_id] : T(0);
    for (int r = 0; r < maxn; ++r) {
      const unsigned src = __fns(peers, 0, r + 1);  // r-th member of my peer set, or ~0u
      const T vv = __shfl_sync(FULL, v, src & 31);
      if (lane == leader && src != 0xffffffffu) acc = add_rn(acc, vv);
    }
    if (act && lane == leader) vals_base[slot_id] = acc;
  }
  __syncwarp();
}

}  // namespace b200sp
