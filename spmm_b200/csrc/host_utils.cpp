// host_utils.cpp -- host-side plumbing of the row-panel driver (no device code).
//
// The reference has no multi-GPU path; its single-GPU drivers read whole matrices and hand them
// to cuSPARSE (cupy_cusparse/spgemm_from_txt_alg1.cu:130-195).  The row-panel driver
// (spmm_b200/multi.py) cuts A into panels of equal WORK, which needs the
// intermediate products of every row before anything is on the GPU.  In numpy that is four passes
// over nnz(A) int64 values (70 - 80 ms for the 5.4e6 nonzeros of two C3 panels - more than the
// 34 ms the panels' C takes to cross PCIe); here it is one multithreaded pass.
#include <stdint.h>

#include <algorithm>
#include <atomic>
#include <thread>
#include <vector>

#include "../../include/b200sp.h"

extern "C" int b200sp_host_row_products(int64_t m, int64_t k, const int32_t *hAp, const int32_t *hAj,
                                        const int32_t *hBp, int64_t *out, int nthreads) {
  if (m < 0 || k < 0 || (m && (!hAp || !out)) || (k && !hBp)) return B200SP_ERR_INVALID;
  if (m == 0) return B200SP_OK;
  const int64_t nnz = hAp[m];
  if (hAp[0] != 0 || nnz < 0 || (nnz && !hAj)) return B200SP_ERR_INVALID;
  int nt = nthreads > 0 ? nthreads : (int)std::min<unsigned>(16u, std::max(1u, std::thread::hardware_concurrency()));
  nt = (int)std::max<int64_t>(1, std::min<int64_t>(nt, (nnz + m) / 65536 + 1));  // small inputs: no threads
  std::atomic<int> bad(0);
  auto work = [&](int t) {
    // rows [r0, r1): equal shares of (nonzeros + rows), found on the row pointers
    auto cut = [&](int q) -> int64_t {
      if (q <= 0) return 0;
      if (q >= nt) return m;
      const int64_t target = (nnz + m) * q / nt;
      int64_t lo = 0, hi = m;
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if ((int64_t)hAp[mid] + mid < target) lo = mid + 1;
        else hi = mid;
      }
      return lo;
    };
    const int64_t r0 = cut(t), r1 = cut(t + 1);
    int b = 0;
    for (int64_t i = r0; i < r1; ++i) {
      const int32_t s = hAp[i], e = hAp[i + 1];
      int64_t sum = 0;
      if (s < 0 || e < s || e > nnz) {
        b = 1;
      } else {
        for (int32_t p = s; p < e; ++p) {
          const int32_t c = hAj[p];
          if (c < 0 || c >= k) {
            b = 1;
            continue;
          }
          sum += (int64_t)hBp[c + 1] - hBp[c];
        }
      }
      out[i] = sum;
    }
    if (b) bad.store(1);
  };
  if (nt == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    th.reserve(nt);
    for (int t = 0; t < nt; ++t) th.emplace_back(work, t);
    for (auto &x : th) x.join();
  }
  return bad.load() ? B200SP_ERR_INVALID : B200SP_OK;
}
