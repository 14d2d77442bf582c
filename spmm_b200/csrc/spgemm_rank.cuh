// spgemm_rank.cuh -- numeric pass for the heavy rows of WIDE matrices: compact rank accumulator.
//
// k_num_dense_pieces (spgemm_numeric.cuh) walks ceil(n / W) dense column blocks per row whether or
// not the row touches them, and every (A entry, block) pair with a non-empty segment costs a
// barrier step: at R-MAT scale 22 (n = 4.2 M, 164 blocks) a B row of ~1000 entries is cut into
// ~30-entry segments and the kernel runs at 2-3 % of its roofline (profiles/r01_panel_cost_rmat22.txt:
// time follows segments, not products).  Here the accumulator is COMPACT: slot = rank of the column
// among the columns the row really touches.  One pass holds CAP consecutive ranks, so a row needs
// ceil(nnz(C_i) / CAP) passes instead of ceil(n / W) blocks - one pass for most heavy rows - and a
// B row is applied whole: pieces of up to NT entries, 180-230 products per barrier step on R-MAT
// 16-20 instead of 20-130.
//
// Per row (one CTA, rows handed out by a global ticket):
//  1. structure: all warps walk the B rows and set one bit per touched column in the CTA's slice
//     of a global (L2-resident) table of {bits, prefix} words - the same walk the symbolic pass did;
//  2. scan: exclusive prefix of the popcounts -> `prefix`; the sorted column indices go straight to
//     C.indices (the rank of a column IS its position in the output row);
//  3. passes: for rank range [p CAP, (p+1) CAP): zero the accumulators, stage the B rows of the A
//     entries in storage order (for multi-pass rows only the part whose columns fall into the pass:
//     B rows are sorted, two binary searches), then one PIECE of one B row per barrier step:
//     slot = prefix[col >> 5] + popc(bits & below) - p CAP, acc[slot] += a * b (separately rounded:
//     the ordered-sum contract of common.cuh; columns inside a B row are distinct, the barrier
//     orders the rows).  The (column, value) loads run 2 batches ahead and the table lookups one
//     batch ahead of the piece being applied, so the barrier chain waits on shared memory only;
//  4. C.data[row, p CAP ...] = alpha * acc - a plain coalesced copy; 5. the bits are cleared.
// Requires canonical B (verified by the symbolic pass).
#pragma once
#include "spgemm_numeric.cuh"

namespace b200sp {

constexpr int RANK_PB = 4;                 // pieces per pipeline batch
constexpr int RANK_PAD = 3 * RANK_PB;      // zero-length staging entries after the last piece

__device__ __forceinline__ uint2 ld_entry(const uint2 *p) {  // L2 (the table is written by this CTA's atomics)
  uint2 v;
  asm volatile("ld.global.cg.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}

// first index in [lo, hi) with Bj[idx] >= key (Bj sorted inside a row)
__device__ __forceinline__ int lower_bound_col(const int *__restrict__ Bj, int lo, int hi, int key) {
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(Bj + mid) < key) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// The pieces staged in s_seg[0 .. cnt) applied in order.  Three-stage software pipeline over
// batches of RANK_PB pieces: (column, value) loads of batch k+2 and table lookups of batch k+1 are
// issued before batch k is applied.
template <typename T>
__device__ __forceinline__ void rank_apply_staged(T *__restrict__ acc, const uint2 *__restrict__ ent,
                                                  const DenseSeg<T> *s_seg, int cnt, int tid, int rank0,
                                                  const int *__restrict__ Bj, const T *__restrict__ Bx) {
  constexpr int PB = RANK_PB;
  int j2[PB], j1[PB], slot0[PB];
  T b2[PB], b1[PB], p0[PB];
  uint2 e1[PB];
  bool a2[PB], a1[PB], a0[PB];
  int l0[PB];
#define RANK_LOAD(Q)                                  \
  _Pragma("unroll") for (int s = 0; s < PB; ++s) {    \
    const DenseSeg<T> sg = s_seg[(Q) + s];            \
    a2[s] = tid < sg.len;                             \
    j2[s] = 0;                                        \
    b2[s] = T(0);                                     \
    if (a2[s]) {                                      \
      j2[s] = ld_inplace(Bj + sg.start + tid);        \
      b2[s] = ld_inplace(Bx + sg.start + tid);        \
    }                                                 \
  }
#define RANK_SHIFT_AND_LOOKUP()                       \
  _Pragma("unroll") for (int s = 0; s < PB; ++s) {    \
    j1[s] = j2[s];                                    \
    b1[s] = b2[s];                                    \
    a1[s] = a2[s];                                    \
    e1[s] = make_uint2(0u, 0u);                       \
    if (a1[s]) e1[s] = ld_entry(ent + (j1[s] >> 5));  \
  }
#define RANK_RESOLVE(Q)                                                                   \
  _Pragma("unroll") for (int s = 0; s < PB; ++s) {                                        \
    const DenseSeg<T> sg = s_seg[(Q) + s];                                                \
    a0[s] = a1[s];                                                                        \
    l0[s] = sg.len;                                                                       \
    slot0[s] = (int)e1[s].y + __popc(e1[s].x & ((1u << (j1[s] & 31)) - 1u)) - rank0;      \
    p0[s] = mul_rn(sg.a, b1[s]);                                                          \
  }
  // prologue: batch 0 loaded and looked up, batch 1 loaded
  RANK_LOAD(0)
  RANK_SHIFT_AND_LOOKUP()
  RANK_LOAD(PB)
  for (int q = 0; q < cnt; q += PB) {
    RANK_RESOLVE(q)              // batch q: slots from the lookups issued one iteration ago
    RANK_SHIFT_AND_LOOKUP()      // batch q + PB: lookups go out (its columns arrived during the last apply)
    RANK_LOAD(q + 2 * PB)        // batch q + 2 PB: (column, value) loads go out
#pragma unroll
    for (int s = 0; s < PB; ++s) {
      if (a0[s]) acc[slot0[s]] = add_rn(acc[slot0[s]], p0[s]);
      if (l0[s] != 0) __syncthreads();  // CTA-uniform; only the padding after the last piece is empty
    }
  }
#undef RANK_LOAD
#undef RANK_SHIFT_AND_LOOKUP
#undef RANK_RESOLVE
}

template <typename T, int NT>
__global__ void __launch_bounds__(NT, 2)
k_num_rank_cta(const int *__restrict__ rows, int nrows, int *__restrict__ ticket, int n, int CAP,
               const int *__restrict__ Ap, const int *__restrict__ Aj, const T *__restrict__ Ax,
               const int *__restrict__ Bp, const int *__restrict__ Bj, const T *__restrict__ Bx,
               const int *__restrict__ Cp, int *__restrict__ Cj, T *__restrict__ Cx, T alpha,
               uint2 *__restrict__ gws) {
  constexpr int NW = NT / 32, SEGCAP = 2 * NT;
  extern __shared__ __align__(16) unsigned char smem_raw[];  // CAP accumulators
  __shared__ DenseSeg<T> s_seg[SEGCAP + RANK_PAD];
  __shared__ int s_scan[NW];
  __shared__ int s_wsum[NW];
  __shared__ int s_ticket;
  T *acc = reinterpret_cast<T *>(smem_raw);
  const int tid = threadIdx.x, lane = lane_id(), w = tid >> 5;
  const int nwords = (n + 31) >> 5;
  uint2 *ent = gws + (size_t)blockIdx.x * nwords;  // {bits, exclusive prefix}; bits are zero on entry to a row
  // contiguous word range of warp w for the scans, a multiple of 32 words
  const int wseg = (((nwords + NW - 1) / NW) + 31) & ~31;
  const int w_lo = min(w * wseg, nwords), w_hi = min(w_lo + wseg, nwords);
  while (true) {
    if (tid == 0) s_ticket = atomicAdd(ticket, 1);
    __syncthreads();
    const int ridx = s_ticket;
    if (ridx >= nrows) break;
    const int row = rows[ridx];
    const int a_begin = __ldg(Ap + row), a_end = __ldg(Ap + row + 1);
    const int out = __ldg(Cp + row), c = __ldg(Cp + row + 1) - out;
    // ---- 1. structure: one bit per touched column (test first: hot columns are hit many times)
    walk_b_columns(a_begin, a_end, w, NW, lane, Aj, Bp, Bj, [&](int j) {
      unsigned *word = &ent[j >> 5].x;
      const unsigned bit = 1u << (j & 31);
      if (!(__ldcg(word) & bit)) atomicOr(word, bit);
    });
    __syncthreads();
    // ---- 2. exclusive prefix of the popcounts; sorted columns -> C.indices
    int mine = 0;
    for (int i = w_lo + lane; i < w_hi; i += 32) mine += __popc(ld_entry(ent + i).x);
    mine = warp_sum(mine);
    if (lane == 0) s_wsum[w] = mine;
    __syncthreads();
    int base = 0;
#pragma unroll
    for (int q = 0; q < NW; ++q)
      if (q < w) base += s_wsum[q];
    for (int i0 = w_lo; i0 < w_hi; i0 += 32) {
      const int i = i0 + lane;
      unsigned bits = i < w_hi ? ld_entry(ent + i).x : 0u;
      const int cnt = __popc(bits);
      int inc = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += u;
      }
      if (bits) {
        int pos = base + inc - cnt;
        ent[i].y = (unsigned)pos;
        int *cj = Cj + out + pos;
        const int colbase = i << 5;
        while (bits) {
          const int bit = __ffs(bits) - 1;
          bits &= bits - 1;
          *cj++ = colbase + bit;
        }
      }
      base += __shfl_sync(FULL, inc, 31);
    }
    __syncthreads();  // prefixes and C.indices of this row are complete
    // ---- 3. / 4. passes over rank ranges of CAP accumulators
    const int npass = (c + CAP - 1) / CAP;
    for (int p = 0; p < npass; ++p) {
      const int rank0 = p * CAP, cap_p = min(CAP, c - rank0);
      // column window of the pass (multi-pass rows only): [clo, chi)
      const bool windowed = npass > 1;
      const int clo = p > 0 ? __ldcg(Cj + out + rank0) : 0;
      const int chi = p + 1 < npass ? __ldcg(Cj + out + rank0 + CAP) : INT_MAX;
      for (int s = tid; s < cap_p; s += NT) acc[s] = T(0);
      __syncthreads();
      for (int abase = a_begin; abase < a_end; abase += NT) {
        // stage: thread t takes A nonzero abase + t; its B row (or the part inside the window) is cut
        // into pieces of <= NT entries; slots from an exclusive scan of the piece counts (A order)
        const int t = abase + tid;
        int start = 0, len = 0;
        T av = T(0);
        if (t < a_end) {
          const int kk = __ldg(Aj + t);
          start = __ldg(Bp + kk);
          int end = __ldg(Bp + kk + 1);
          if (windowed && end > start) {
            if (p > 0) start = lower_bound_col(Bj, start, end, clo);
            if (p + 1 < npass) end = lower_bound_col(Bj, start, end, chi);
          }
          len = end - start;
          av = __ldg(Ax + t);
        }
        const int pieces = (len + NT - 1) / NT;
        int inc = pieces;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int u = __shfl_up_sync(FULL, inc, o);
          if (lane >= o) inc += u;
        }
        if (lane == 31) s_scan[w] = inc;
        __syncthreads();
        int off = inc - pieces, total = 0;
#pragma unroll
        for (int q = 0; q < NW; ++q) {
          const int v = s_scan[q];
          if (q < w) off += v;
          total += v;
        }
        for (int r0 = 0; r0 < total; r0 += SEGCAP) {
          const int cnt = min(SEGCAP, total - r0);
          const int p_lo = max(0, r0 - off), p_hi = min(pieces, r0 + SEGCAP - off);
          for (int pc = p_lo; pc < p_hi; ++pc) {
            DenseSeg<T> sg;
            sg.start = start + pc * NT;
            sg.len = min(NT, len - pc * NT);
            sg.a = av;
            s_seg[off + pc - r0] = sg;
          }
          if (tid < RANK_PAD) {
            DenseSeg<T> z;
            z.start = 0;
            z.len = 0;
            z.a = T(0);
            s_seg[cnt + tid] = z;
          }
          __syncthreads();
          rank_apply_staged<T>(acc, ent, s_seg, cnt, tid, rank0, Bj, Bx);
          __syncthreads();
        }
        if (total == 0) __syncthreads();  // s_scan is rewritten by the next chunk
      }
      // values of the pass: a plain copy of the compact accumulator
      T *cx = Cx + out + rank0;
      if (alpha == T(1)) {
        for (int s = tid; s < cap_p; s += NT) cx[s] = acc[s];
      } else {
        for (int s = tid; s < cap_p; s += NT) cx[s] = mul_rn(alpha, acc[s]);
      }
      __syncthreads();
    }
    // ---- 5. clear the bits for the next row
    for (int i = w_lo + lane; i < w_hi; i += 32)
      if (ld_entry(ent + i).x) ent[i].x = 0u;
    __syncthreads();
  }
}

}  // namespace b200sp
