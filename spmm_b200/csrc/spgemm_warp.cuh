// spgemm_warp.cuh -- heavy rows of wide matrices, one WARP per (row, narrow column block).
//
// k_num_dense_pieces (spgemm_numeric.cuh) gives a whole CTA to a (row, 25 600-column block): every
// B-row segment costs a CTA-wide barrier step of 8 warps x ~33 instructions however short it is.
// On power-law inputs most segments are short (R-MAT scale 18: the non-empty segments average 81
// entries, 62 % of them hold fewer than 32), so five warps of eight idle through most steps:
// 16.3 G warp instructions for 2.9 G products, issue slots 36 % busy, 3 barrier stalls per issue
// (profiles/r02_ncu_full_rmat18_pieces_v17.txt).
//
// Here the unit of work is a warp with its own accumulator of WW columns (10 KB) and its own
// touched-column bitmap: 16 independent warps per SM, no CTA barrier anywhere.  A task is a
// (row, block) pair; tasks are handed out by a global ticket, a few at a time (the same block of a
// few consecutive rows).  Per task the warp
//   1. visits the row's A entries 128 at a time, lane-parallel: the segment of B row k inside the
//      block is [bounds[k][b], bounds[k][b+1]) (the table of k_block_bounds at WW granularity),
//      empty segments cost nothing further;
//   2. queues the non-empty segments in A order as steps of at most 32 entries in its own shared
//      memory, and applies them one after the other - lane l adds a * b_l to acc[col_l], the
//      columns inside a step are distinct, steps follow each other in program order (__syncwarp),
//      so every column's sum is in A-row storage order; the (column, value) loads run WB_D steps
//      ahead of the step being applied;
//   3. counts its touched columns, learns where its entries start inside the row from the counts of
//      the row's earlier blocks (decoupled look-back over per-task status words, as in the single
//      pass of k_num_dense_cta), and emits them in column order.
#pragma once
#include "common.cuh"
#include "spgemm_numeric.cuh"

namespace b200sp {

constexpr int WB_NT = 128;      // threads per CTA: 4 independent warps
constexpr int WB_D = 8;         // steps whose loads are in flight
constexpr int WB_Q = 160;       // steps staged per flush (per warp)
constexpr int WB_QPAD = 2 * WB_D;
constexpr int WB_CHUNK = 4;     // tasks per ticket draw

struct __align__(8) WbStep {
  int start, cnt;
};

#define WB_FLAG_AGG (1u << 30)
#define WB_FLAG_PREFIX (2u << 30)
#define WB_VALUE_MASK ((1u << 30) - 1u)

// Entries of the row that lie in blocks [0, b): look-back over the status words of the row's earlier
// tasks (task id t = ridx * nblk + b; only tasks of the same row are looked at).
__device__ __forceinline__ int wb_lookback(volatile unsigned *status, long long t, int b, int own, int lane) {
  if (b == 0) {
    if (lane == 0) status[t] = WB_FLAG_PREFIX | (unsigned)own;
    return 0;
  }
  if (lane == 0) status[t] = WB_FLAG_AGG | (unsigned)own;
  int excl = 0;
  for (int j = b - 1;; j -= 32) {  // lane l looks at block j - l of this row
    const int bb = j - lane;
    unsigned s;
    int fp;
    for (;;) {
      s = bb >= 0 ? status[t - b + bb] : WB_FLAG_PREFIX;
      const unsigned flag = s >> 30;
      const unsigned pref = __ballot_sync(FULL, flag == 2u), inval = __ballot_sync(FULL, flag == 0u);
      fp = pref ? __ffs(pref) - 1 : 32;
      const unsigned upto = fp < 32 ? (2u << fp) - 1u : FULL;
      if ((inval & upto) == 0) break;
      __nanosleep(40);
    }
    int v = lane <= fp ? (int)(s & WB_VALUE_MASK) : 0;
    v = warp_sum(v);
    excl += v;
    if (fp < 32) break;
  }
  if (lane == 0) status[t] = WB_FLAG_PREFIX | (unsigned)(excl + own);
  return excl;
}

// The steps staged in qs / qa [0 .. cnt) applied in order by one warp (see dense_apply_staged: the
// same software pipeline at warp scope, __syncwarp instead of the CTA barrier).
template <typename T, int D>
__device__ __forceinline__ void wb_apply_staged(T *__restrict__ acc, unsigned *__restrict__ bm, const WbStep *qs,
                                                const T *qa, int cnt, int lane, int c0, const int *__restrict__ Bj,
                                                const T *__restrict__ Bx, int last) {
  int j[D];
  T b[D];
  WbStep nx = qs[0];
#define WB_FETCH(S, Q)                           \
  {                                              \
    const int idx_ = min(nx.start + lane, last); \
    nx = qs[(Q) + 1];                            \
    j[S] = ld_inplace(Bj + idx_);                \
    b[S] = ld_inplace(Bx + idx_);                \
  }
#pragma unroll
  for (int s = 0; s < D; ++s) WB_FETCH(s, s)
  for (int q = 0; q < cnt; q += D) {
#pragma unroll
    for (int s = 0; s < D; ++s) {
      const int len = qs[q + s].cnt;  // warp-uniform (broadcast read)
      const T a = qa[q + s];
      const bool act = lane < len;
      const int lc = j[s] - c0;
      T cur = T(0);
      if (act) cur = acc[lc];
      const T prod = add_rn(mul_rn(a, b[s]), T(0));  // -0.0 -> +0.0
      WB_FETCH(s, q + D + s)
      if (act) {
        acc[lc] = add_rn(cur, prod);
        if (NegZero<T>::is(cur)) atomicOr(bm + (lc >> 5), 1u << (lc & 31));
      }
      __syncwarp();
    }
  }
#undef WB_FETCH
}

template <typename T>
__global__ void __launch_bounds__(WB_NT, 4)
k_num_warp_blocks(const int *__restrict__ rows, int nrows, int *__restrict__ ticket, int n, int WW, int nblk,
                  const int *__restrict__ Ap, const int *__restrict__ Aj, const T *__restrict__ Ax,
                  const int *__restrict__ bounds, const int *__restrict__ Bj, const T *__restrict__ Bx, int nnzB,
                  const int *__restrict__ Cp, int *__restrict__ Cj, T *__restrict__ Cx, T alpha,
                  unsigned *__restrict__ status) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = lane_id(), w = threadIdx.x >> 5;
  const int nw = WW >> 5;  // bitmap words per block; WW is a multiple of 32
  const size_t per_warp = (size_t)WW * sizeof(T) + (size_t)(WB_Q + WB_QPAD) * (sizeof(WbStep) + sizeof(T)) + (size_t)nw * 4;
  unsigned char *mine = smem_raw + (size_t)w * ((per_warp + 15) & ~(size_t)15);
  T *acc = reinterpret_cast<T *>(mine);
  WbStep *qs = reinterpret_cast<WbStep *>(mine + (size_t)WW * sizeof(T));
  T *qa = reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(qs) + (size_t)(WB_Q + WB_QPAD) * sizeof(WbStep));
  unsigned *bm = reinterpret_cast<unsigned *>(reinterpret_cast<unsigned char *>(qa) + (size_t)(WB_Q + WB_QPAD) * sizeof(T));
  const T UNT = NegZero<T>::value();
  for (int c = lane; c < WW; c += 32) acc[c] = UNT;
  for (int c = lane; c < nw; c += 32) bm[c] = 0u;
  __syncwarp();
  // Task order: rows in sets of WB_CHUNK; inside a set block after block, and inside a block the set's rows.
  // The WB_CHUNK consecutive tasks one ticket hands to a warp are therefore the SAME block of DIFFERENT rows:
  // a task waits for the counts of its row's earlier blocks, and with consecutive blocks of one row in one
  // warp (first version) every row was processed serially along that chain (R-MAT scale 18: 1140 ms).
  const int nsets = (nrows + WB_CHUNK - 1) / WB_CHUNK;
  const long long ntasks = (long long)nsets * WB_CHUNK * nblk;
  const int stride = nblk + 1;
  for (;;) {
    int t0 = 0;
    if (lane == 0) t0 = atomicAdd(ticket, WB_CHUNK);
    t0 = __shfl_sync(FULL, t0, 0);
    if (t0 >= ntasks) break;
    const int t1 = (int)min((long long)t0 + WB_CHUNK, ntasks);
    for (int t = t0; t < t1; ++t) {
      const int set = t / (WB_CHUNK * nblk), rem = t - set * (WB_CHUNK * nblk);
      const int b = rem / WB_CHUNK, ridx = set * WB_CHUNK + (rem - b * WB_CHUNK);
      if (ridx >= nrows) continue;  // the last set may be short
      const int row = __ldg(rows + ridx);
      const int a_begin = __ldg(Ap + row), a_end = __ldg(Ap + row + 1);
      const int c0 = b * WW;
      // ---- accumulate: 128 A entries per round (4 per lane, A order = (u, lane)) ----
      for (int base = a_begin; base < a_end; base += 128) {
        int st[4], ln[4], ns[4], off[4];
        T av[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int e = base + u * 32 + lane;
          st[u] = 0, ln[u] = 0, av[u] = T(0);
          if (e < a_end) {
            const int kk = __ldg(Aj + e);
            const int *bd = bounds + (size_t)kk * stride + b;
            st[u] = __ldg(bd);
            ln[u] = __ldg(bd + 1) - st[u];
            av[u] = __ldg(Ax + e);
          }
        }
        int total = 0;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          ns[u] = (ln[u] + 31) >> 5;
          int inc = ns[u];
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += v;
          }
          off[u] = total + inc - ns[u];
          total += __shfl_sync(FULL, inc, 31);
        }
        for (int r0 = 0; r0 < total; r0 += WB_Q) {
          const int cnt = min(WB_Q, total - r0);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int p_lo = max(0, r0 - off[u]), p_hi = min(ns[u], r0 + WB_Q - off[u]);
            for (int pc = p_lo; pc < p_hi; ++pc) {
              WbStep sg;
              sg.start = st[u] + (pc << 5);
              sg.cnt = min(32, ln[u] - (pc << 5));
              qs[off[u] + pc - r0] = sg;
              qa[off[u] + pc - r0] = av[u];
            }
          }
          if (lane < WB_QPAD) {
            WbStep z;
            z.start = 0, z.cnt = 0;
            qs[cnt + lane] = z;
            qa[cnt + lane] = T(0);
          }
          __syncwarp();
          wb_apply_staged<T, WB_D>(acc, bm, qs, qa, cnt, lane, c0, Bj, Bx, nnzB - 1);
          __syncwarp();
        }
      }
      // ---- count, offset inside the row, ordered emit ----
      int mine_cnt = 0;
      for (int i = lane; i < nw; i += 32) mine_cnt += __popc(bm[i]);
      const int own = warp_sum(mine_cnt);
      const int excl = wb_lookback(status, (long long)ridx * nblk + b, b, own, lane);
      if (own == 0) continue;  // warp-uniform
      int out = __ldg(Cp + row) + excl;
      for (int i0 = 0; i0 < nw; i0 += 32) {
        const int i = i0 + lane;
        unsigned word = i < nw ? bm[i] : 0u;
        const int c = __popc(word);
        int inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int v = __shfl_up_sync(FULL, inc, o);
          if (lane >= o) inc += v;
        }
        const int tot = __shfl_sync(FULL, inc, 31);
        if (tot == 0) continue;  // warp-uniform
        if (word) bm[i] = 0u;
        int *cj = Cj + (out + inc - c);
        T *cx = Cx + (out + inc - c);
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
        out += tot;
      }
      __syncwarp();
    }
  }
}

}  // namespace b200sp
