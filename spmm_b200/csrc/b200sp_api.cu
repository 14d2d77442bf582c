// b200sp_api.cu -- C ABI of libb200sp.so (see include/b200sp.h) and host orchestration.
//
// Host side of the hot path: owns the workspace arena (with a high-water mark, the analogue of
// the CuPy MemoryPool the reference samples: SpGEMM_alg_comparison/profiler.py:82-143), plans
// the bins, launches the sm_100a kernels and performs the two host synchronisations the
// algorithm needs (bin sizes after the product count; nnz(C) after the symbolic pass - the
// latter is the same sync the reference has at cupyx/cusparse.py:2120-2127).
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b200sp.h"
#include "spgemm_numeric.cuh"
#include "spgemm_setup.cuh"
#include "spgemm_symbolic.cuh"
#include "spgemm_rank.cuh"
#include "spgemm_warp.cuh"
#include "spmv.cuh"
#include "csr_utils.cuh"
#include "convert.cuh"

using namespace b200sp;

namespace {

constexpr size_t SMEM_MAX = 227 * 1024;      // opt-in dynamic shared memory per CTA on sm_100
constexpr size_t SMEM_BITMAP_MAX = 200 * 1024;  // budget for an n-bit bitmap in shared memory
constexpr size_t SMEM_DENSE_MAX = 208 * 1024;   // budget for an n-column dense accumulator
constexpr size_t SMEM_BLOCKED_MAX = 100 * 1024;  // accumulator budget of k_num_dense_pieces: two CTAs per SM (measured best)
constexpr int DENSE_BLOCKED_MAX_BLOCKS = 384;    // most column blocks it may need (R-MAT scale 22 in fp64: 328)
constexpr int PIECES_NT = 256;                   // its CTA size = the longest piece of a B row applied per barrier
constexpr int RANKBM_MAX_N = 32768;
constexpr size_t SMEM_RANK_ACC = 96 * 1024;     // compact accumulator of k_num_rank_cta: two CTAs of 256 threads per SM
constexpr int RANK_NT = 256;             // widest matrix for the bitmap-rank output of the warp bins

struct Buffer {  // grow-only device buffer
  void *ptr = nullptr;
  size_t size = 0;
};

enum SymKind { SYM_GROUP8, SYM_WARP, SYM_CTA_HASH, SYM_CTA_BITMAP };
enum NumKind { NUM_GROUP8, NUM_WARP, NUM_CTA_HASH, NUM_GLOBAL, NUM_DENSE, NUM_DENSE_BLOCKED, NUM_RANK, NUM_WARP_BLOCKS };

struct BinPlan {
  int kind;
  int S;  // table slots (hash kinds)
};

struct Plan {
  bool valid = false;
  int m = 0, k = 0, n = 0;
  int64_t nnzA = 0, nnzB = 0;
  const int *Ap = nullptr, *Aj = nullptr, *Bp = nullptr, *Bj = nullptr;
  int *Cp = nullptr;
  int alg = 0;
  int num_chunks = 1;
  BinLimits sym_limits{}, num_limits{};
  BinPlan sym_bins[MAX_BINS]{}, num_bins[MAX_BINS]{};
  int sym_count[MAX_BINS]{}, num_count[MAX_BINS]{};
  int *rows_sym = nullptr, *rows_num = nullptr;  // device, inside ws
  int64_t num_products = 0, nnzC = 0, max_row_products = 0;
  bool heavy_ordered = false;  // the last numeric bin's row list is ordered by decreasing work and its bucket histogram is in ctr->sym_count
  bool stage1 = false;      // row products / bound / symbolic bins of the current arguments are in ws (estimate ran)
  bool single_pass = false; // the estimate chose the single-pass kernel
  int64_t nnz_bound = 0;    // sum_i min(products_i, n)
  bool b_canonical = true;  // B rows verified sorted (needed by the column-blocked dense kernel)
  int bounds_nblk = 0, bounds_W = 0;  // geometry of the per-B-row block-boundary table now in ws3 (0 = none)
};

}  // namespace

struct b200sp_context {
  int device = 0;
  int num_sms = 148;  // queried at create (cudaDevAttrMultiProcessorCount); sizes every persistent grid
  cudaStream_t stream = nullptr;
  void *(*alloc_fn)(size_t, void *) = nullptr;
  void (*free_fn)(void *, void *) = nullptr;
  void *alloc_ctx = nullptr;
  Buffer ws;       // plan workspace (row lists, counters, scan tiles)
  Buffer ws2;      // global bitmap workspace of the heavy-row kernels
  Buffer ws3;      // row ticket + per-B-row column-block boundaries of k_num_dense_pieces
  Buffer io[10];   // device staging of the *_host entry points
  void *pinned_out[3] = {nullptr, nullptr, nullptr};  // host staging of C (handle owned)
  size_t pinned_out_size[3] = {0, 0, 0};
  Counters *h_counters = nullptr;  // pinned
  int64_t bytes_live = 0, bytes_peak = 0;
  int64_t launches = 0;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaStream_t copy_stream = nullptr;  // second DMA queue of the *_host entry points (ingest pipeline)
  cudaEvent_t ev_copy = nullptr, ev_sym = nullptr;
  cudaEvent_t ev_slot[2] = {nullptr, nullptr}, ev_chunk = nullptr;  // chunked (ALG3) host entry: output double buffer
  float ms_symbolic = 0.f, ms_numeric = 0.f;
  bool numeric_timed = false;
  Plan plan;
  // SpMV kernel choice: longest row of the last matrix seen (keyed by indptr pointer and sizes; a
  // stale hit only costs speed, never correctness)
  const int *spmv_key_ap = nullptr;
  int spmv_key_m = -1;
  int64_t spmv_key_nnz = -1;
  int spmv_max_row = 0;
  // launch geometry of the single-pass kernel, cached per (value size, CTA size, shared memory)
  int sp_key_t = 0, sp_key_nt = 0, sp_per_sm = 0;
  size_t sp_key_smem = 0;
  std::string err;
};

namespace {

int fail(b200sp_context *h, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (h) h->err = buf;
  return code;
}

#define CUDA_TRY(h, expr)                                                                    \
  do {                                                                                       \
    cudaError_t e_ = (expr);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      int code_ = (e_ == cudaErrorMemoryAllocation) ? B200SP_ERR_ALLOC : B200SP_ERR_CUDA;    \
      return fail(h, code_, "CUDA error %s:%d: %s (%s)", __FILE__, __LINE__,                 \
                  cudaGetErrorString(e_), #expr);                                            \
    }                                                                                        \
  } while (0)

#define TRY(expr)              \
  do {                         \
    int rc_ = (expr);          \
    if (rc_ != B200SP_OK) return rc_; \
  } while (0)

int dev_alloc(b200sp_context *h, void **p, size_t bytes) {
  *p = nullptr;
  if (bytes == 0) return B200SP_OK;
  if (h->alloc_fn) {
    *p = h->alloc_fn(bytes, h->alloc_ctx);
    if (!*p) return fail(h, B200SP_ERR_ALLOC, "caller allocator returned NULL for %zu bytes", bytes);
  } else {
    cudaError_t e = cudaMalloc(p, bytes);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(h, B200SP_ERR_ALLOC, "out of device memory allocating %zu bytes: %s", bytes,
                  cudaGetErrorString(e));
    }
  }
  h->bytes_live += (int64_t)bytes;
  h->bytes_peak = std::max(h->bytes_peak, h->bytes_live);
  return B200SP_OK;
}

void dev_free(b200sp_context *h, void *p, size_t bytes) {
  if (!p) return;
  if (h->free_fn) h->free_fn(p, h->alloc_ctx);
  else cudaFree(p);
  h->bytes_live -= (int64_t)bytes;
}

// Grow-only: contents are NOT preserved across a grow.
int ensure(b200sp_context *h, Buffer &b, size_t bytes) {
  if (bytes <= b.size) return B200SP_OK;
  if (b.ptr) {
    cudaStreamSynchronize(h->stream);
    dev_free(h, b.ptr, b.size);
    b.ptr = nullptr;
    b.size = 0;
  }
  const size_t want = (bytes + 255) & ~(size_t)255;
  TRY(dev_alloc(h, &b.ptr, want));
  b.size = want;
  return B200SP_OK;
}

void release(b200sp_context *h, Buffer &b) {
  if (b.ptr) dev_free(h, b.ptr, b.size);
  b.ptr = nullptr;
  b.size = 0;
}

template <typename K>
int set_smem(b200sp_context *h, K kernel, size_t bytes) {
  if (bytes > 48 * 1024) {
    CUDA_TRY(h, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  }
  return B200SP_OK;
}

int check_launch(b200sp_context *h, const char *what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(h, B200SP_ERR_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(e));
  h->launches++;
  return B200SP_OK;
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr size_t align_up_c(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace

// =============================================================================== handle ====

extern "C" int b200sp_version(void) { return B200SP_VERSION; }

extern "C" const char *b200sp_status_string(int status) {
  switch (status) {
    case B200SP_OK: return "ok";
    case B200SP_ERR_INVALID: return "invalid argument";
    case B200SP_ERR_CUDA: return "CUDA error";
    case B200SP_ERR_ALLOC: return "out of memory";
    case B200SP_ERR_OVERFLOW: return "int32 overflow";
    case B200SP_ERR_STATE: return "bad call sequence";
    case B200SP_ERR_IO: return "I/O error";
    default: return "unknown status";
  }
}

extern "C" int b200sp_create(b200sp_handle_t *handle, int device) {
  if (!handle) return B200SP_ERR_INVALID;
  *handle = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return B200SP_ERR_CUDA;  // no CPU fallback: the product needs a CUDA device
  }
  b200sp_context *h = new (std::nothrow) b200sp_context();
  if (!h) return B200SP_ERR_ALLOC;
  if (device < 0) {
    if (cudaGetDevice(&device) != cudaSuccess) device = 0;
  }
  h->device = device;
  DeviceGuard guard(device);  // the caller's current device is restored on return
  int cur = -1;
  if (cudaGetDevice(&cur) != cudaSuccess || cur != device ||
      cudaMallocHost((void **)&h->h_counters, sizeof(Counters)) != cudaSuccess) {
    cudaGetLastError();
    delete h;
    return B200SP_ERR_CUDA;
  }
  if (cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || h->num_sms <= 0) {
    cudaGetLastError();
    h->num_sms = 148;
  }
  for (auto &e : h->ev) cudaEventCreate(&e);
  cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking);
  cudaEventCreateWithFlags(&h->ev_copy, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&h->ev_sym, cudaEventDisableTiming);
  for (auto &e : h->ev_slot) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&h->ev_chunk, cudaEventDisableTiming);
  *handle = h;
  return B200SP_OK;
}

extern "C" int b200sp_release_workspace(b200sp_handle_t h) {
  if (!h) return B200SP_ERR_INVALID;
  DeviceGuard guard(h->device);
  cudaStreamSynchronize(h->stream);
  release(h, h->ws);
  release(h, h->ws2);
  release(h, h->ws3);
  for (auto &b : h->io) release(h, b);
  h->plan.valid = false;
  return B200SP_OK;
}

extern "C" int b200sp_destroy(b200sp_handle_t h) {
  if (!h) return B200SP_OK;
  DeviceGuard guard(h->device);
  b200sp_release_workspace(h);
  for (int i = 0; i < 3; ++i)
    if (h->pinned_out[i]) cudaFreeHost(h->pinned_out[i]);
  if (h->h_counters) cudaFreeHost(h->h_counters);
  for (auto &e : h->ev)
    if (e) cudaEventDestroy(e);
  if (h->ev_copy) cudaEventDestroy(h->ev_copy);
  if (h->ev_sym) cudaEventDestroy(h->ev_sym);
  for (auto &e : h->ev_slot)
    if (e) cudaEventDestroy(e);
  if (h->ev_chunk) cudaEventDestroy(h->ev_chunk);
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  delete h;
  return B200SP_OK;
}

extern "C" int b200sp_set_stream(b200sp_handle_t h, void *cuda_stream) {
  if (!h) return B200SP_ERR_INVALID;
  if (h->stream != (cudaStream_t)cuda_stream) {
    // the workspaces (and the plan's row lists) are reused by the next call: work still queued on
    // the old stream must be done before anything on the new one may overwrite or free them
    DeviceGuard guard(h->device);
    cudaStreamSynchronize(h->stream);
    h->stream = (cudaStream_t)cuda_stream;
  }
  return B200SP_OK;
}

extern "C" int b200sp_set_allocator(b200sp_handle_t h, void *(*alloc_fn)(size_t, void *),
                                    void (*free_fn)(void *, void *), void *ctx) {
  if (!h || ((alloc_fn == nullptr) != (free_fn == nullptr))) return B200SP_ERR_INVALID;
  b200sp_release_workspace(h);  // buffers must be freed by the allocator that made them
  h->alloc_fn = alloc_fn;
  h->free_fn = free_fn;
  h->alloc_ctx = ctx;
  return B200SP_OK;
}

extern "C" const char *b200sp_last_error(b200sp_handle_t h) { return h ? h->err.c_str() : "null handle"; }

extern "C" int b200sp_get_stats(b200sp_handle_t h, b200sp_stats *s) {
  if (!h || !s) return B200SP_ERR_INVALID;
  memset(s, 0, sizeof *s);
  if (h->numeric_timed) {
    cudaEventSynchronize(h->ev[3]);
    cudaEventElapsedTime(&h->ms_numeric, h->ev[2], h->ev[3]);
  }
  const Plan &p = h->plan;
  s->num_products = p.num_products;
  s->nnz_c = p.nnzC;
  s->workspace_bytes = h->bytes_live;
  s->peak_workspace_bytes = h->bytes_peak;
  s->max_row_products = p.max_row_products;
  s->num_chunks = p.num_chunks;
  s->num_bins = p.sym_limits.nbins;
  for (int b = 0; b < MAX_BINS; ++b) {
    s->sym_bin_rows[b] = p.sym_count[b];
    s->num_bin_rows[b] = p.num_count[b];
  }
  s->kernel_launches = h->launches;
  s->ms_symbolic = h->ms_symbolic;
  s->ms_numeric = h->ms_numeric;
  return B200SP_OK;
}

extern "C" int b200sp_reset_stats(b200sp_handle_t h) {
  if (!h) return B200SP_ERR_INVALID;
  h->bytes_peak = h->bytes_live;
  h->launches = 0;
  return B200SP_OK;
}

// =============================================================================== SpGEMM ====

namespace {

void plan_bins(Plan &p) {
  const size_t bitmap_bytes = (size_t)((p.n + 31) / 32) * 4;
  // symbolic bins, by the upper bound ub = min(products, n) of nnz(C_i)
  const int sl[6] = {32, 128, 512, 4096, 16384, INT_MAX};
  p.sym_limits.nbins = 6;
  for (int b = 0; b < 6; ++b) p.sym_limits.limit[b] = sl[b];
  p.sym_bins[0] = {SYM_GROUP8, 64};
  p.sym_bins[1] = {SYM_WARP, 256};
  p.sym_bins[2] = {SYM_WARP, 1024};
  p.sym_bins[3] = {bitmap_bytes <= 8192 * 4 ? SYM_CTA_BITMAP : SYM_CTA_HASH, 8192};
  p.sym_bins[4] = {bitmap_bytes <= 32768 * 4 ? SYM_CTA_BITMAP : SYM_CTA_HASH, 32768};
  p.sym_bins[5] = {SYM_CTA_BITMAP, 0};
  // numeric bins, by the exact nnz(C_i)
  const int nl[6] = {32, 128, 512, 6144, 12288, INT_MAX};
  p.num_limits.nbins = 6;
  for (int b = 0; b < 6; ++b) p.num_limits.limit[b] = nl[b];
  p.num_bins[0] = {NUM_GROUP8, 64};
  p.num_bins[1] = {NUM_WARP, 256};
  p.num_bins[2] = {NUM_WARP, 1024};
  p.num_bins[3] = {NUM_CTA_HASH, 8192};
  p.num_bins[4] = {NUM_CTA_HASH, 16384};
  p.num_bins[5] = {NUM_GLOBAL, 0};
  if ((size_t)p.n * 4 <= SMEM_DENSE_MAX) {
    // few columns: rows above 512 entries use the dense shared-memory accumulator (no sort)
    const int dl[4] = {32, 128, 512, INT_MAX};
    p.num_limits.nbins = 4;
    for (int b = 0; b < 4; ++b) p.num_limits.limit[b] = dl[b];
    p.num_bins[3] = {NUM_DENSE, 0};
  } else {
    // wide matrix.  Rows up to t entries keep the 8 K-slot shared-memory hash (the 16 K-slot table
    // leaves one 8-warp CTA per SM: R-MAT scale 20 runs 10 % faster without it).  Above that:
    //   rank    - compact rank accumulator (k_num_rank_cta): passes = ceil(nnz(C_i) / CAP), B rows applied whole
    //   blocked - dense column blocks (k_num_dense_pieces): ceil(n / W) blocks per row
    // B200SP_HEAVY = rank | blocked | mixed (rank up to n / B200SP_HEAVY_DENSE_DIV entries, blocked for the
    // denser rows); B200SP_HEAVY_MIN overrides t.
    const char *mode = getenv("B200SP_HEAVY");
    const char *tmin = getenv("B200SP_HEAVY_MIN");
    const char *ddiv = getenv("B200SP_HEAVY_DENSE_DIV");
    const bool blocked_ok = (size_t)p.n * 8 <= (size_t)DENSE_BLOCKED_MAX_BLOCKS * SMEM_BLOCKED_MAX;
    // default: blocked.  Measured on B200 (profiles/r02_heavy_modes_v16.jsonl, whole A @ A incl. symbolic):
    // R-MAT scale 16 / 18 / 20: blocked 7.3 / 56 / 698 ms, rank (96 KB) 9.8 / 73 / 927 ms, mixed (n/8) 8.4 / 68 / 848 ms
    const bool want_rank = mode && strcmp(mode, "rank") == 0;
    const bool want_warp = mode && strcmp(mode, "warp") == 0;  // k_num_warp_blocks (falls back to blocked when its table would not fit)
    const bool want_blocked = blocked_ok && !want_rank && !want_warp && !(mode && strcmp(mode, "mixed") == 0);
    const bool want_mixed = blocked_ok && mode && strcmp(mode, "mixed") == 0;
    int t_dense = std::max(512, std::min(6144, p.n / 128));
    if (tmin && atoi(tmin) >= 512) t_dense = std::min(6144, atoi(tmin));
    int nb = 3;
    if (t_dense > 512) {
      p.num_limits.limit[nb] = t_dense;
      p.num_bins[nb++] = {NUM_CTA_HASH, 8192};
    }
    if (want_warp) {
      p.num_limits.limit[nb] = INT_MAX;
      p.num_bins[nb++] = {NUM_WARP_BLOCKS, 0};
    } else if (want_blocked) {
      p.num_limits.limit[nb] = INT_MAX;
      p.num_bins[nb++] = {NUM_DENSE_BLOCKED, 0};
    } else if (want_mixed) {
      const int div = ddiv && atoi(ddiv) > 0 ? atoi(ddiv) : 8;
      const int t_rank = std::max(t_dense + 1, p.n / div);
      p.num_limits.limit[nb] = t_rank;
      p.num_bins[nb++] = {NUM_RANK, 0};
      p.num_limits.limit[nb] = INT_MAX;
      p.num_bins[nb++] = {NUM_DENSE_BLOCKED, 0};
    } else {
      p.num_limits.limit[nb] = INT_MAX;
      p.num_bins[nb++] = {NUM_RANK, 0};
    }
    p.num_limits.nbins = nb;
  }
}

int persistent_grid(b200sp_context *h, int nrows, size_t smem_bytes) {
  int per_sm = smem_bytes ? (int)std::min<size_t>(8, SMEM_MAX / std::max<size_t>(smem_bytes, 1)) : 8;
  per_sm = std::max(per_sm, 1);
  return std::max(1, std::min(nrows, h->num_sms * per_sm));
}

int launch_symbolic_bin(b200sp_context *h, int b, const int *rows, int cnt, int *ticket = nullptr) {
  Plan &p = h->plan;
  const BinPlan &bp = p.sym_bins[b];
  cudaStream_t st = h->stream;
  switch (bp.kind) {
    case SYM_GROUP8: {
      const size_t smem = 16 * (size_t)bp.S * 4;
      k_sym_hash_group<8><<<(cnt + 15) / 16, 128, smem, st>>>(rows, cnt, bp.S, p.Ap, p.Aj, p.Bp, p.Bj, p.Cp);
      return check_launch(h, "k_sym_hash_group<8>");
    }
    case SYM_WARP: {
      const size_t smem = 4 * (size_t)bp.S * 4;
      k_sym_hash_group<32><<<(cnt + 3) / 4, 128, smem, st>>>(rows, cnt, bp.S, p.Ap, p.Aj, p.Bp, p.Bj, p.Cp);
      return check_launch(h, "k_sym_hash_group<32>");
    }
    case SYM_CTA_HASH: {
      const size_t smem = (size_t)bp.S * 4;
      TRY(set_smem(h, k_sym_hash_cta, smem));
      k_sym_hash_cta<<<persistent_grid(h, cnt, smem), 256, smem, st>>>(rows, cnt, bp.S, p.Ap, p.Aj, p.Bp, p.Bj, p.Cp);
      return check_launch(h, "k_sym_hash_cta");
    }
    case SYM_CTA_BITMAP: {
      const size_t bytes = (size_t)((p.n + 31) / 32) * 4;
      const size_t fbytes = (size_t)((p.n + 3) / 4) * 4;  // one byte flag per column
      if (fbytes <= SMEM_BITMAP_MAX) {
        TRY(set_smem(h, k_sym_flags_cta, fbytes));
        k_sym_flags_cta<<<persistent_grid(h, cnt, fbytes + 64), 256, fbytes, st>>>(rows, cnt, p.n, p.Ap, p.Aj, p.Bp, p.Bj, p.Cp, ticket);
        return check_launch(h, "k_sym_flags_cta");
      }
      if (bytes <= SMEM_BITMAP_MAX) {
        TRY(set_smem(h, k_sym_bitmap_cta<true>, bytes));
        // a warp walks one B row at a time with nothing in flight behind it: the kernel lives on
        // resident warps, so a bitmap that leaves room for one or two CTAs per SM gets bigger CTAs
        const int per_sm = (int)std::max<size_t>(1, SMEM_MAX / (bytes + 64));
        const int nt = per_sm >= 4 ? 256 : per_sm >= 2 ? 512 : 1024;
        k_sym_bitmap_cta<true><<<persistent_grid(h, cnt, bytes + 64), nt, bytes, st>>>(
            rows, cnt, p.n, p.Ap, p.Aj, p.Bp, p.Bj, p.Cp, nullptr, ticket);
        return check_launch(h, "k_sym_bitmap_cta<smem>");
      }
      int grid = std::max(1, std::min(cnt, 2 * h->num_sms / p.num_chunks));
      TRY(ensure(h, h->ws2, (size_t)grid * bytes));
      CUDA_TRY(h, cudaMemsetAsync(h->ws2.ptr, 0, (size_t)grid * bytes, st));
      k_sym_bitmap_cta<false><<<grid, 1024, 0, st>>>(rows, cnt, p.n, p.Ap, p.Aj, p.Bp, p.Bj, p.Cp,
                                                    (unsigned *)h->ws2.ptr, ticket);
      return check_launch(h, "k_sym_bitmap_cta<global>");
    }
  }
  return fail(h, B200SP_ERR_INVALID, "bad symbolic bin kind");
}

// The packed, bank-dealt copy of B the dense kernel reads (spgemm_numeric.cuh: k_pack_b), rebuilt by every
// value pass (the values may be new) into ws3: one kernel over B, ~10 us on C3.
template <typename T>
int pack_b(b200sp_context *h, const T *Bx, const PBEnt<T> **pb, int *pb_last) {
  Plan &p = h->plan;
  const size_t entries = (((size_t)p.nnzB + 31) & ~(size_t)31) + 32 * (size_t)p.k;
  if (entries > (size_t)INT32_MAX) return fail(h, B200SP_ERR_OVERFLOW, "packed copy of B does not fit int32 offsets");
  TRY(ensure(h, h->ws3, entries * sizeof(PBEnt<T>)));
  p.bounds_nblk = p.bounds_W = 0;  // ws3 no longer holds a block-boundary table
  const int grid = std::max(1, std::min((p.k + 7) / 8, h->num_sms * 8));
  k_pack_b<T><<<grid, 256, 0, h->stream>>>(p.k, p.Bp, p.Bj, Bx, (PBEnt<T> *)h->ws3.ptr);
  TRY(check_launch(h, "k_pack_b"));
  *pb = (const PBEnt<T> *)h->ws3.ptr;
  *pb_last = (int)entries - 1;
  return B200SP_OK;
}

// prefetch depth of k_num_dense_cta: 12 when the accumulator alone limits the SM to three CTAs (see DENSE_DEPTH_MAX)
inline bool dense_deep(size_t acc_bytes) { return acc_bytes * 4 > SMEM_MAX - 4 * 6144; }

template <typename T>
int launch_numeric_bin(b200sp_context *h, int b, const int *rows, int cnt, T alpha, const T *Ax,
                       const T *Bx, int *Cj, T *Cx) {
  Plan &p = h->plan;
  const BinPlan &bp = p.num_bins[b];
  cudaStream_t st = h->stream;
  switch (bp.kind) {
    case NUM_GROUP8: {
      const size_t smem = 16 * (size_t)bp.S * (sizeof(T) + 4);
      k_num_hash_group<8, T, false><<<(cnt + 15) / 16, 128, smem, st>>>(rows, cnt, bp.S, p.n, p.Ap, p.Aj, Ax, p.Bp, p.Bj, Bx, p.Cp, Cj, Cx, alpha);
      return check_launch(h, "k_num_hash_group<8>");
    }
    case NUM_WARP: {
      if (p.n <= RANKBM_MAX_N) {  // few columns: bitmap rank instead of the in-row sort
        const size_t smem = 4 * ((size_t)bp.S * (sizeof(T) + 4) + 2 * (size_t)((p.n + 31) / 32) * 4);
        TRY(set_smem(h, k_num_hash_group<32, T, true>, smem));
        k_num_hash_group<32, T, true><<<(cnt + 3) / 4, 128, smem, st>>>(rows, cnt, bp.S, p.n, p.Ap, p.Aj, Ax, p.Bp, p.Bj, Bx, p.Cp, Cj, Cx, alpha);
        return check_launch(h, "k_num_hash_group<32,rank>");
      }
      const size_t smem = 4 * (size_t)bp.S * (sizeof(T) + 4);
      TRY(set_smem(h, k_num_hash_group<32, T, false>, smem));
      k_num_hash_group<32, T, false><<<(cnt + 3) / 4, 128, smem, st>>>(rows, cnt, bp.S, p.n, p.Ap, p.Aj, Ax, p.Bp, p.Bj, Bx, p.Cp, Cj, Cx, alpha);
      return check_launch(h, "k_num_hash_group<32>");
    }
    case NUM_CTA_HASH: {
      const size_t smem = (size_t)bp.S * (sizeof(T) + 4);
      TRY(set_smem(h, k_num_hash_cta<T>, smem));
      k_num_hash_cta<T><<<persistent_grid(h, cnt, smem + 2304), 256, smem, st>>>(rows, cnt, bp.S, p.Ap, p.Aj, Ax, p.Bp, p.Bj, Bx, p.Cp, Cj, Cx, alpha);
      return check_launch(h, "k_num_hash_cta");
    }
    case NUM_DENSE:
      if ((size_t)p.n * sizeof(T) <= SMEM_DENSE_MAX) {
        const size_t smem = (((size_t)p.n + 127) & ~(size_t)127) * sizeof(T);  // padded to whole 128-column groups
        const double avgB = p.k ? (double)p.nnzB / (double)p.k : 0.0;
        const int nt = std::max(64, std::min(DENSE_MAX_NT, (int)((avgB + 31) / 32) * 32));
        const PBEnt<T> *pb = nullptr;
        int pb_last = 0;
        TRY(pack_b<T>(h, Bx, &pb, &pb_last));
        if (dense_deep(smem)) {
          TRY(set_smem(h, k_num_dense_cta<T, false, 12>, smem));
          k_num_dense_cta<T, false, 12><<<persistent_grid(h, cnt, smem + 7424), nt, smem, st>>>(rows, cnt, p.n, p.Ap, p.Aj, Ax, p.Bp, pb, pb_last, p.Cp, Cj, Cx, alpha, SinglePassArgs{});
        } else {
          TRY(set_smem(h, k_num_dense_cta<T, false, 8>, smem));
          k_num_dense_cta<T, false, 8><<<persistent_grid(h, cnt, smem + 7424), nt, smem, st>>>(rows, cnt, p.n, p.Ap, p.Aj, Ax, p.Bp, pb, pb_last, p.Cp, Cj, Cx, alpha, SinglePassArgs{});
        }
        return check_launch(h, "k_num_dense_cta");
      }
      // fp64 with n between the fp32 and fp64 budgets: column blocks
    case NUM_DENSE_BLOCKED:
      if (p.b_canonical && (size_t)p.n * sizeof(T) <= (size_t)DENSE_BLOCKED_MAX_BLOCKS * SMEM_BLOCKED_MAX) {
        // two CTAs of 256 threads per SM; B200SP_BLOCKED_KB / B200SP_BLOCKED_NT = experiment switches (accumulator
        // kilobytes per CTA, 128 or 256 threads per CTA)
        size_t budget = SMEM_BLOCKED_MAX;
        int nt = PIECES_NT;
        if (const char *e = getenv("B200SP_BLOCKED_KB")) {
          const long v = atol(e);
          if (v >= 8 && v <= 200) budget = (size_t)v * 1024;
        }
        if (const char *e = getenv("B200SP_BLOCKED_NT")) nt = atoi(e) == 128 ? 128 : PIECES_NT;
        const int nblk = (int)(((size_t)p.n * sizeof(T) + budget - 1) / budget);
        const int W = (((p.n + nblk - 1) / nblk) + 127) & ~127;  // whole 128-column groups
        const size_t smem = (size_t)W * sizeof(T) + (size_t)(W / 32) * 4;  // accumulators + touched-column bitmap
        if (p.bounds_nblk != nblk || p.bounds_W != W) {  // first numeric call of this plan / value type
          const size_t total = (size_t)p.k * (nblk + 1);
          TRY(ensure(h, h->ws3, 256 + total * sizeof(int)));  // [row ticket | pad | table]
          const int bgrid = (int)std::max<size_t>(1, std::min<size_t>((total + 255) / 256, (size_t)h->num_sms * 16));
          k_block_bounds<<<bgrid, 256, 0, st>>>(p.k, nblk, W, p.Bp, p.Bj, (int *)h->ws3.ptr + 64);
          TRY(check_launch(h, "k_block_bounds"));
          p.bounds_nblk = nblk;
          p.bounds_W = W;
        }
        // B200SP_WINDOW_TURN = windows per hand-off of the short-segment path (1, 4 or 8), B200SP_FLAT_BELOW =
        // products per piece below which a chunk takes it (experiment switches; defaults measured on B200)
        // (profiles/r02_heavy_modes_window_turns.jsonl: numeric pass at R-MAT scale 16 / 18 / 20 with 1, 4, 8 windows
        // per turn 5.06 / 48.3 / 585, 4.67 / 48.4 / 624, 5.35 / 50.7 / 712 ms; raising FLAT_BELOW to 64 - 128 loses everywhere:
        // the window path is bound by locating and loading its products, not by the hand-off)
        int wpt = 1, flat_below = PIECES_FLAT_BELOW;
        if (const char *e = getenv("B200SP_WINDOW_TURN")) wpt = atoi(e);
        if (const char *e = getenv("B200SP_FLAT_BELOW")) flat_below = std::max(0, atoi(e));
        auto kern = nt == 128 ? k_num_dense_pieces<T, 128, 1>
                    : wpt >= 8 ? k_num_dense_pieces<T, PIECES_NT, 8> : wpt >= 4 ? k_num_dense_pieces<T, PIECES_NT, 4> : k_num_dense_pieces<T, PIECES_NT, 1>;
        TRY(set_smem(h, kern, smem));
        int per_sm = 1;
        CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, nt, smem));
        const int grid = std::max(1, std::min(cnt, h->num_sms * std::max(1, per_sm)));
        CUDA_TRY(h, cudaMemsetAsync(h->ws3.ptr, 0, 2 * sizeof(int), st));  // rows are handed out by ticket ([1]: the hub pre-pass)
        const int *hub_hist = nullptr, *hub_offs = nullptr;
        int grid_main = grid;
        // ... when the largest row alone exceeds a CTA's fair share of the products, i.e. is the critical path (the
        // pre-pass walks the hub rows' structure once more: measured at N = 1 it costs 13 % / 4 % of the numeric pass
        // at R-MAT scale 16 / 18, where no row is that large, and gains 1 % at scale 20, where the top rows are)
        if (p.heavy_ordered && b == p.num_limits.nbins - 1 && !getenv("B200SP_NO_HUB_SPLIT") &&
            p.max_row_products * (int64_t)(h->num_sms * std::max(1, per_sm)) >= p.num_products) {
          // the heaviest rows (top two work buckets of the ordered list) are cut into one task per column block
          Counters *ctr = (Counters *)h->ws.ptr;
          TRY(ensure(h, h->ws2, (size_t)HUB_MAX * (nblk + 1) * sizeof(int)));
          int *counts = (int *)h->ws2.ptr;
          const size_t hsmem = (size_t)(W / 32) * 4;
          k_hub_counts<<<h->num_sms * 4, 256, hsmem, st>>>(rows, cnt, &ctr->sym_count[0], (int *)h->ws3.ptr + 1, W, nblk, p.Ap, p.Aj,
                                                          (const int *)h->ws3.ptr + 64, p.Bj, counts);
          TRY(check_launch(h, "k_hub_counts"));
          k_hub_offsets<<<(HUB_MAX * 32 + 255) / 256, 256, 0, st>>>(&ctr->sym_count[0], cnt, nblk, counts);
          TRY(check_launch(h, "k_hub_offsets"));
          hub_hist = &ctr->sym_count[0];
          hub_offs = counts;
          grid_main = h->num_sms * std::max(1, per_sm);  // the task count (rows + hub blocks) is only known on the device
        }
        kern<<<grid_main, nt, smem, st>>>(rows, cnt, (int *)h->ws3.ptr, p.n, W, nblk, p.Ap, p.Aj, Ax, (const int *)h->ws3.ptr + 64, p.Bj, Bx, (int)p.nnzB, p.Cp, Cj, Cx, alpha, flat_below, hub_hist, hub_offs);
        return check_launch(h, "k_num_dense_pieces");
      }
      // B not canonical (or absurdly wide): the global-memory kernel takes any input
      goto global_fallback;
    case NUM_WARP_BLOCKS: {
      int WW = sizeof(T) == 4 ? 2560 : 1280;  // 10 KB of accumulator per warp: 16 warps per SM
      if (const char *e = getenv("B200SP_WARP_WW")) {
        const int v = atoi(e) & ~31;
        if (v >= 256 && v <= 16384) WW = v;
      }
      const int nblk = (p.n + WW - 1) / WW;
      const size_t table = (size_t)p.k * (nblk + 1) * sizeof(int);
      if (!p.b_canonical || table > ((size_t)3 << 30) || ((long long)cnt + WB_CHUNK) * nblk > (long long)INT32_MAX - 2 * WB_CHUNK) {
        // the boundary table or the task ids would not fit: the CTA-per-block kernel takes these rows
        BinPlan saved = p.num_bins[b];
        p.num_bins[b].kind = NUM_DENSE_BLOCKED;
        const int rc = launch_numeric_bin<T>(h, b, rows, cnt, alpha, Ax, Bx, Cj, Cx);
        p.num_bins[b] = saved;
        return rc;
      }
      if (p.bounds_nblk != nblk || p.bounds_W != WW) {
        const size_t total = (size_t)p.k * (nblk + 1);
        TRY(ensure(h, h->ws3, 256 + total * sizeof(int)));  // [row ticket | pad | table]
        const int bgrid = (int)std::max<size_t>(1, std::min<size_t>((total + 255) / 256, (size_t)h->num_sms * 16));
        k_block_bounds<<<bgrid, 256, 0, st>>>(p.k, nblk, WW, p.Bp, p.Bj, (int *)h->ws3.ptr + 64);
        TRY(check_launch(h, "k_block_bounds"));
        p.bounds_nblk = nblk;
        p.bounds_W = WW;
      }
      const size_t per_warp = ((size_t)WW * sizeof(T) + (size_t)(WB_Q + WB_QPAD) * (sizeof(WbStep) + sizeof(T)) + (size_t)(WW / 32) * 4 + 15) & ~(size_t)15;
      const size_t smem = per_warp * (WB_NT / 32);
      TRY(set_smem(h, k_num_warp_blocks<T>, smem));
      int per_sm = 1;
      CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_num_warp_blocks<T>, WB_NT, smem));
      const long long ntasks = (long long)cnt * nblk;  // status words; the ticket runs over whole sets of WB_CHUNK rows
      const long long want = (ntasks + WB_CHUNK * (WB_NT / 32) - 1) / (WB_CHUNK * (WB_NT / 32));
      const int grid = (int)std::max<long long>(1, std::min<long long>(want, (long long)h->num_sms * std::max(1, per_sm)));
      TRY(ensure(h, h->ws2, (size_t)ntasks * sizeof(unsigned)));  // per-task status words of the look-back
      CUDA_TRY(h, cudaMemsetAsync(h->ws2.ptr, 0, (size_t)ntasks * sizeof(unsigned), st));
      CUDA_TRY(h, cudaMemsetAsync(h->ws3.ptr, 0, sizeof(int), st));  // the task ticket
      k_num_warp_blocks<T><<<grid, WB_NT, smem, st>>>(rows, cnt, (int *)h->ws3.ptr, p.n, WW, nblk, p.Ap, p.Aj, Ax, (const int *)h->ws3.ptr + 64, p.Bj, Bx, (int)p.nnzB, p.Cp, Cj, Cx, alpha, (unsigned *)h->ws2.ptr);
      return check_launch(h, "k_num_warp_blocks");
    }
    case NUM_RANK: {
      size_t acc_bytes = SMEM_RANK_ACC;
      if (const char *kb = getenv("B200SP_RANK_ACC_KB")) {
        const long v = atol(kb);
        if (v >= 8 && v <= 200) acc_bytes = (size_t)v * 1024;
      }
      const int CAP = (int)(acc_bytes / sizeof(T)) & ~31;
      const size_t smem = (size_t)CAP * sizeof(T);
      TRY(set_smem(h, k_num_rank_cta<T, RANK_NT>, smem));
      int per_sm = 1;
      CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_num_rank_cta<T, RANK_NT>, RANK_NT, smem));
      const int grid = std::max(1, std::min(cnt, h->num_sms * std::max(1, per_sm) / (p.num_chunks > 1 ? 2 : 1)));
      const size_t nwords = (size_t)((p.n + 31) / 32);
      TRY(ensure(h, h->ws2, (size_t)grid * nwords * sizeof(uint2)));  // {bits, prefix} per 32 columns and CTA
      TRY(ensure(h, h->ws3, 256));
      CUDA_TRY(h, cudaMemsetAsync(h->ws2.ptr, 0, (size_t)grid * nwords * sizeof(uint2), st));
      CUDA_TRY(h, cudaMemsetAsync(h->ws3.ptr, 0, sizeof(int), st));  // rows are handed out by ticket
      k_num_rank_cta<T, RANK_NT><<<grid, RANK_NT, smem, st>>>(rows, cnt, (int *)h->ws3.ptr, p.n, CAP, p.Ap, p.Aj, Ax, p.Bp, p.Bj, Bx, p.Cp, Cj, Cx, alpha, (uint2 *)h->ws2.ptr);
      return check_launch(h, "k_num_rank_cta");
    }
    global_fallback:
    case NUM_GLOBAL: {
      const size_t bytes = (size_t)((p.n + 31) / 32) * 8;  // bitmap + prefix
      int grid = std::max(1, std::min(cnt, 2 * h->num_sms / p.num_chunks));
      TRY(ensure(h, h->ws2, (size_t)grid * bytes));
      CUDA_TRY(h, cudaMemsetAsync(h->ws2.ptr, 0, (size_t)grid * bytes, st));
      k_num_bitmap_global<T><<<grid, 256, 0, st>>>(rows, cnt, p.n, p.Ap, p.Aj, Ax, p.Bp, p.Bj, Bx, p.Cp, Cj, Cx, alpha, (unsigned *)h->ws2.ptr);
      return check_launch(h, "k_num_bitmap_global");
    }
  }
  return fail(h, B200SP_ERR_INVALID, "bad numeric bin kind");
}

template <typename T>
int run_numeric(b200sp_context *h, const void *alpha, const void *Ax, const void *Bx, int *Cj, void *Cx) {
  Plan &p = h->plan;
  const T a = alpha ? *(const T *)alpha : T(1);
  int off = 0;
  for (int b = 0; b < p.num_limits.nbins; ++b) {
    const int cnt = p.num_count[b];
    // ALG2 / ALG3 on device-resident operands: C is caller-owned and allocated whole, so the only
    // memory left to bound is the spill workspace of the global-bitmap kernels - their grid (and
    // with it the CTA-private workspace) shrinks by num_chunks; every other bin runs as in ALG1
    // (round 1 cut every bin's row list into num_chunks launches: time for nothing).
    if (cnt > 0) TRY(launch_numeric_bin<T>(h, b, p.rows_num + off, cnt, a, (const T *)Ax, (const T *)Bx, Cj, (T *)Cx));
    off += cnt;
  }
  return B200SP_OK;
}

}  // namespace

namespace {

struct WsLayout {
  size_t o_ub, o_rs, o_rn, o_ts, o_st, total;
  int ntiles;
};

// plan workspace: counters | rowub | rows_sym | rows_num | scan tile sums | single-pass row status
WsLayout ws_layout(int m) {
  WsLayout L;
  L.ntiles = (m + SCAN_TILE - 1) / SCAN_TILE;
  L.o_ub = align_up(sizeof(Counters), 256);
  L.o_rs = L.o_ub + align_up((size_t)m * 4, 256);
  L.o_rn = L.o_rs + align_up((size_t)m * 4, 256);
  L.o_ts = L.o_rn + align_up((size_t)m * 4, 256);
  L.o_st = L.o_ts + align_up((size_t)L.ntiles * 8, 256);
  L.total = L.o_st + (size_t)m * 8;
  return L;
}

// Stage 1 of the structure pass: products per row, their total, the bound sum_i min(products_i, n)
// of nnz(C), the symbolic bins, the canonical check of B.  Ends with host sync #1.
int spgemm_stage1(b200sp_context *h, int64_t m, int64_t k, int64_t n, int64_t nnzA, int64_t nnzB,
                  const int32_t *dAp, const int32_t *dAj, const int32_t *dBp, const int32_t *dBj, int alg,
                  float chunk_fraction, int32_t *dCp, bool enqueue_only = false) {
  Plan &p = h->plan;
  p.valid = false;
  p.stage1 = false;
  p.single_pass = false;
  p.heavy_ordered = false;
  if (m < 0 || k < 0 || n < 0 || nnzA < 0 || nnzB < 0 || m >= INT32_MAX || k >= INT32_MAX || n >= INT32_MAX ||
      nnzA > INT32_MAX || nnzB > INT32_MAX)
    return fail(h, B200SP_ERR_INVALID, "bad shape m=%lld k=%lld n=%lld nnzA=%lld nnzB=%lld", (long long)m,
                (long long)k, (long long)n, (long long)nnzA, (long long)nnzB);
  if (!dAp || !dBp || !dCp || (nnzA && !dAj) || (nnzB && !dBj))
    return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  const bool chunked = (alg == B200SP_ALG2 || alg == B200SP_ALG3);
  if (chunked && !(chunk_fraction > 0.f && chunk_fraction <= 1.f))
    return fail(h, B200SP_ERR_INVALID, "chunk_fraction must be in (0,1], got %f", chunk_fraction);
  cudaStream_t st = h->stream;
  p.m = (int)m; p.k = (int)k; p.n = (int)n; p.nnzA = nnzA; p.nnzB = nnzB;
  p.Ap = dAp; p.Aj = dAj; p.Bp = dBp; p.Bj = dBj; p.Cp = dCp;
  p.alg = alg;
  p.num_chunks = chunked ? (int)std::min<double>(64.0, std::ceil(1.0 / (double)chunk_fraction - 1e-6)) : 1;
  p.num_products = 0; p.nnzC = 0; p.max_row_products = 0; p.nnz_bound = 0;
  p.bounds_nblk = p.bounds_W = 0;
  memset(p.sym_count, 0, sizeof p.sym_count);
  memset(p.num_count, 0, sizeof p.num_count);
  plan_bins(p);
  h->numeric_timed = false;
  if (m == 0) {
    CUDA_TRY(h, cudaMemsetAsync(dCp, 0, sizeof(int), st));
    p.stage1 = true;
    return B200SP_OK;
  }
  const WsLayout L = ws_layout(p.m);
  TRY(ensure(h, h->ws, L.total));
  char *base = (char *)h->ws.ptr;
  Counters *ctr = (Counters *)base;
  int *rowub = (int *)(base + L.o_ub);
  int *rowprod = (int *)(base + L.o_st);  // products per row; shares the single-pass status area (used after it is dead)
  p.rows_sym = (int *)(base + L.o_rs);
  p.rows_num = (int *)(base + L.o_rn);

  cudaEventRecord(h->ev[0], st);
  CUDA_TRY(h, cudaMemsetAsync(ctr, 0, sizeof(Counters), st));
  {
    const double avg = (double)nnzA / (double)m;
    const int G = avg <= 2 ? 2 : avg <= 6 ? 4 : avg <= 12 ? 8 : avg <= 24 ? 16 : 32;
    const int grid = std::max(1, std::min((int)((m + 256 / G - 1) / (256 / G)), h->num_sms * 8));
    switch (G) {
      case 2: k_row_products<2><<<grid, 256, 0, st>>>(p.m, p.n, dAp, dAj, dBp, rowub, dCp, p.sym_limits, ctr, rowprod); break;
      case 4: k_row_products<4><<<grid, 256, 0, st>>>(p.m, p.n, dAp, dAj, dBp, rowub, dCp, p.sym_limits, ctr, rowprod); break;
      case 8: k_row_products<8><<<grid, 256, 0, st>>>(p.m, p.n, dAp, dAj, dBp, rowub, dCp, p.sym_limits, ctr, rowprod); break;
      case 16: k_row_products<16><<<grid, 256, 0, st>>>(p.m, p.n, dAp, dAj, dBp, rowub, dCp, p.sym_limits, ctr, rowprod); break;
      default: k_row_products<32><<<grid, 256, 0, st>>>(p.m, p.n, dAp, dAj, dBp, rowub, dCp, p.sym_limits, ctr, rowprod); break;
    }
    TRY(check_launch(h, "k_row_products"));
  }
  // Every numeric kernel applies one B row (or a piece of it) per ordered step and relies on the
  // columns inside a B row being distinct; the column-blocked kernels also on their being sorted.
  // One O(nnzB) pass verifies it (the precondition asserted at cupyx/cusparse.py:2030-2031); the
  // flag comes back with the counters of sync #1.
  if (nnzB > 0) {
    const int cgrid = std::max(1, std::min((int)((k + 7) / 8), h->num_sms * 8));
    k_csr_check_canonical<<<cgrid, 256, 0, st>>>(p.k, p.n, dBp, dBj, &ctr->b_bad);
    TRY(check_launch(h, "k_csr_check_canonical"));
  }
  if (enqueue_only) {  // b200sp_spgemm_single_pass_direct: no row lists, and the counters are read once, at its end
    cudaEventRecord(h->ev[1], st);
    return B200SP_OK;
  }
  const int sgrid = std::max(1, std::min((p.m + 255) / 256, h->num_sms * 8));
  k_scatter_rows<false><<<sgrid, 256, 0, st>>>(p.m, rowub, p.sym_limits, ctr, p.rows_sym);
  TRY(check_launch(h, "k_scatter_rows<sym>"));
  cudaEventRecord(h->ev[1], st);  // (re-recorded at the end of stage 2)
  // sync #1: product total + symbolic bin sizes
  CUDA_TRY(h, cudaMemcpyAsync(h->h_counters, ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaStreamSynchronize(st));
  if (h->h_counters->overflow)
    return fail(h, B200SP_ERR_OVERFLOW, "a row of A@B has more than 2^31-1 intermediate products");
  if (h->h_counters->b_bad)
    return fail(h, B200SP_ERR_INVALID,
                "B is not canonical (a row has unsorted, duplicate or out-of-range column indices): "
                "canonicalise it first (b200sp_csr_sum_duplicates / b200sp_csr_sort_indices)");
  p.num_products = (int64_t)h->h_counters->total_products;
  p.max_row_products = h->h_counters->max_row_products;
  p.nnz_bound = (int64_t)h->h_counters->total_ub;
  for (int b = 0; b < p.sym_limits.nbins; ++b) p.sym_count[b] = h->h_counters->sym_count[b];
  p.b_canonical = true;
  p.stage1 = true;
  return B200SP_OK;
}

// Stage 2: exact row sizes per symbolic bin, scan to C.indptr, numeric bins.  Ends with host sync #2.
int spgemm_stage2(b200sp_context *h, int64_t *nnzC) {
  Plan &p = h->plan;
  cudaStream_t st = h->stream;
  int32_t *dCp = p.Cp;
  const WsLayout L = ws_layout(p.m);
  char *base = (char *)h->ws.ptr;
  Counters *ctr = (Counters *)base;
  long long *tile_sums = (long long *)(base + L.o_ts);
  const int ntiles = L.ntiles;
  const int sgrid = std::max(1, std::min((p.m + 255) / 256, h->num_sms * 8));
  int off = 0;
  for (int b = 0; b < p.sym_limits.nbins; ++b) {
    const int cnt = p.sym_count[b];
    int *ticket = nullptr;
    // the last bin (the rows with the most products, no upper limit) of a bitmap / flag kernel: rows by
    // decreasing work, handed out by ticket (see k_work_order)
    if (b == p.sym_limits.nbins - 1 && p.sym_bins[b].kind == SYM_CTA_BITMAP && cnt > 2 * h->num_sms &&
        p.max_row_products >= 64 && !getenv("B200SP_NO_HEAVY_FIRST")) {
      const int *rowprod = (const int *)(base + L.o_st);
      CUDA_TRY(h, cudaMemsetAsync(&ctr->sym_count[0], 0, 32 * sizeof(int), st));  // host copies were taken at sync #1
      CUDA_TRY(h, cudaMemsetAsync(&ctr->ticket, 0, sizeof(int), st));
      int top_log2 = 0;
      while ((p.max_row_products >> (top_log2 + 1)) > 0) ++top_log2;
      const int og = std::max(1, std::min((cnt + 255) / 256, h->num_sms * 8));
      k_work_hist<<<og, 256, 0, st>>>(p.rows_sym + off, cnt, rowprod, top_log2, &ctr->sym_count[0]);
      TRY(check_launch(h, "k_work_hist"));
      k_work_order<<<og, 256, 0, st>>>(p.rows_sym + off, cnt, rowprod, top_log2, &ctr->sym_count[0], &ctr->sym_cursor[0], p.rows_num);
      TRY(check_launch(h, "k_work_order"));
      CUDA_TRY(h, cudaMemcpyAsync(p.rows_sym + off, p.rows_num, (size_t)cnt * sizeof(int), cudaMemcpyDeviceToDevice, st));
      ticket = &ctr->ticket;
    }
    if (cnt > 0) TRY(launch_symbolic_bin(h, b, p.rows_sym + off, cnt, ticket));
    off += cnt;
  }
  // counts -> indptr
  const bool small_scan = p.m <= 4096;  // measured: one block beats the 3-kernel tiled scan only up to here
  if (small_scan) {  // also fills the numeric-bin histogram
    k_scan_single<<<1, SCAN1_THREADS, 0, st>>>(dCp, p.m, p.num_limits, ctr);
    TRY(check_launch(h, "k_scan_single"));
  } else {
    k_scan_tile_sums<<<ntiles, SCAN_THREADS, 0, st>>>(dCp, p.m, tile_sums);
    TRY(check_launch(h, "k_scan_tile_sums"));
    k_scan_tile_offsets<<<1, SCAN_THREADS, 0, st>>>(tile_sums, ntiles, dCp, p.m, ctr);
    TRY(check_launch(h, "k_scan_tile_offsets"));
    k_scan_apply<<<ntiles, SCAN_THREADS, 0, st>>>(dCp, p.m, tile_sums);
    TRY(check_launch(h, "k_scan_apply"));
  }
  if (!small_scan) {
    k_hist_numeric<<<sgrid, 256, 0, st>>>(p.m, dCp, p.num_limits, ctr);
    TRY(check_launch(h, "k_hist_numeric"));
  }
  k_scatter_rows<true><<<sgrid, 256, 0, st>>>(p.m, dCp, p.num_limits, ctr, p.rows_num);
  TRY(check_launch(h, "k_scatter_rows<num>"));
  cudaEventRecord(h->ev[1], st);
  // sync #2: nnz(C) + numeric bin sizes (the reference's cusparseSpMatGetSize sync)
  CUDA_TRY(h, cudaMemcpyAsync(h->h_counters, ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaStreamSynchronize(st));
  cudaEventElapsedTime(&h->ms_symbolic, h->ev[0], h->ev[1]);
  p.nnzC = (int64_t)h->h_counters->nnz_c;
  for (int b = 0; b < p.num_limits.nbins; ++b) p.num_count[b] = h->h_counters->num_count[b];
  {  // the rows of the last (ticket-scheduled) bin by decreasing work: see k_work_order
    const int lb = p.num_limits.nbins - 1;
    const int kind = p.num_bins[lb].kind, cnt = p.num_count[lb];
    const bool ticketed = kind == NUM_DENSE_BLOCKED || kind == NUM_RANK || kind == NUM_WARP_BLOCKS;
    if (ticketed && cnt > 2 * h->num_sms && p.max_row_products >= 64 && !getenv("B200SP_NO_HEAVY_FIRST")) {
      int off_lb = 0;
      for (int b = 0; b < lb; ++b) off_lb += p.num_count[b];
      const int *rowprod = (const int *)(base + L.o_st);
      static_assert(offsetof(Counters, sym_cursor) == offsetof(Counters, sym_count) + 16 * sizeof(int), "hist | cursor");
      CUDA_TRY(h, cudaMemsetAsync(&ctr->sym_count[0], 0, 32 * sizeof(int), st));  // 16 bucket counts | 16 cursors (both dead by now)
      int top_log2 = 0;
      while ((p.max_row_products >> (top_log2 + 1)) > 0) ++top_log2;
      const int og = std::max(1, std::min((cnt + 255) / 256, h->num_sms * 8));
      k_work_hist<<<og, 256, 0, st>>>(p.rows_num + off_lb, cnt, rowprod, top_log2, &ctr->sym_count[0]);
      TRY(check_launch(h, "k_work_hist"));
      k_work_order<<<og, 256, 0, st>>>(p.rows_num + off_lb, cnt, rowprod, top_log2, &ctr->sym_count[0], &ctr->sym_cursor[0], p.rows_sym);
      TRY(check_launch(h, "k_work_order"));
      CUDA_TRY(h, cudaMemcpyAsync(p.rows_num + off_lb, p.rows_sym, (size_t)cnt * sizeof(int), cudaMemcpyDeviceToDevice, st));
      p.heavy_ordered = true;
    }
  }
  *nnzC = p.nnzC;
  if (p.nnzC > INT32_MAX)
    return fail(h, B200SP_ERR_OVERFLOW, "nnz(C) = %lld does not fit the int32 indptr of CSR", (long long)p.nnzC);
  p.single_pass = false;
  p.valid = true;
  return B200SP_OK;
}

// The single-pass kernel pays one sweep of the n accumulator columns per row whatever the row
// holds, and saves the structure pass (one more walk of the products): it wins where the rows of C
// are well filled.  Measured on B200 (profiles/r02_single_pass_sweep.txt).
bool single_pass_pays(const Plan &p, int dtype) {
  const size_t s = dtype == B200SP_R_64F ? 8 : 4;
  if (const char *e = getenv("B200SP_SINGLE_PASS")) {
    if (strcmp(e, "0") == 0) return false;
  }
  if (p.m == 0 || p.num_chunks > 1 || (size_t)p.n * s > SMEM_DENSE_MAX || p.nnz_bound > INT32_MAX) return false;
  // products per row, in units of n, from which the single pass is chosen.  Measured cross-over on B200
  // (profiles/r02_single_pass_sweep_v17.jsonl): C3 16384^2 level at 0.15 (1.092 vs 1.105 ms), ahead from 0.41 on
  // (d = 1e-2, 1.64: 1.625 -> 1.355 ms); C4 8192^2 ahead from 0.07 on (d = 1e-1, 82: 12.1 -> 8.35 ms)
  double ratio = 0.25;
  if (const char *e = getenv("B200SP_SINGLE_PASS_RATIO")) ratio = atof(e);
  return (double)p.num_products >= ratio * (double)p.m * (double)p.n;
}

}  // namespace

extern "C" int b200sp_spgemm_estimate(b200sp_handle_t h, int dtype, int64_t m, int64_t k, int64_t n, int64_t nnzA,
                                      int64_t nnzB, const int32_t *dAp, const int32_t *dAj, const int32_t *dBp,
                                      const int32_t *dBj, int alg, float chunk_fraction, int32_t *dCp,
                                      int64_t *num_products, int64_t *nnz_bound, int *single_pass) {
  if (!h) return B200SP_ERR_INVALID;
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  DeviceGuard guard(h->device);
  if (num_products) *num_products = 0;
  if (nnz_bound) *nnz_bound = 0;
  if (single_pass) *single_pass = 0;
  TRY(spgemm_stage1(h, m, k, n, nnzA, nnzB, dAp, dAj, dBp, dBj, alg, chunk_fraction, dCp));
  Plan &p = h->plan;
  p.single_pass = single_pass_pays(p, dtype);
  if (num_products) *num_products = p.num_products;
  if (nnz_bound) *nnz_bound = p.nnz_bound;
  if (single_pass) *single_pass = p.single_pass ? 1 : 0;
  return B200SP_OK;
}

extern "C" int b200sp_spgemm_symbolic(b200sp_handle_t h, int64_t m, int64_t k, int64_t n, int64_t nnzA,
                                      int64_t nnzB, const int32_t *dAp, const int32_t *dAj,
                                      const int32_t *dBp, const int32_t *dBj, int alg,
                                      float chunk_fraction, int64_t *num_products, int64_t *nnzC,
                                      int32_t *dCp) {
  if (!h) return B200SP_ERR_INVALID;
  if (!nnzC) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  Plan &p = h->plan;
  DeviceGuard guard(h->device);
  const bool reuse = (alg & B200SP_ALG_REUSE_ESTIMATE) != 0;
  alg &= ~B200SP_ALG_REUSE_ESTIMATE;
  if (num_products) *num_products = 0;
  *nnzC = 0;
  if (reuse) {
    if (!p.stage1 || p.m != m || p.k != k || p.n != n || p.nnzA != nnzA || p.nnzB != nnzB || p.Ap != dAp ||
        p.Aj != dAj || p.Bp != dBp || p.Bj != dBj || p.Cp != dCp || p.alg != alg)
      return fail(h, B200SP_ERR_STATE, "B200SP_ALG_REUSE_ESTIMATE without a matching b200sp_spgemm_estimate");
  } else {
    TRY(spgemm_stage1(h, m, k, n, nnzA, nnzB, dAp, dAj, dBp, dBj, alg, chunk_fraction, dCp));
  }
  p.stage1 = false;  // consumed
  if (num_products) *num_products = p.num_products;
  if (m == 0) {
    p.valid = true;
    return B200SP_OK;
  }
  return spgemm_stage2(h, nnzC);
}

namespace {

template <typename T>
int run_single_pass(b200sp_context *h, const void *alpha, const void *Ax, const void *Bx, int *Cj, void *Cx,
                    bool guarded = false) {
  Plan &p = h->plan;
  const T a = alpha ? *(const T *)alpha : T(1);
  cudaStream_t st = h->stream;
  const WsLayout L = ws_layout(p.m);
  char *base = (char *)h->ws.ptr;
  Counters *ctr = (Counters *)base;
  unsigned long long *status = (unsigned long long *)(base + L.o_st);
  const size_t smem = (((size_t)p.n + 127) & ~(size_t)127) * sizeof(T);
  const double avgB = p.k ? (double)p.nnzB / (double)p.k : 0.0;
  const int nt = std::max(64, std::min(DENSE_MAX_NT, (int)((avgB + 31) / 32) * 32));
  // every CTA must be resident: rows wait for the rows before them.  The attribute call and the occupancy
  // query cost ~10 us of host time between the estimate and the launch: done once per (type, shape) and handle
  const int key_t = (int)sizeof(T);
  if (h->sp_key_t != key_t || h->sp_key_nt != nt || h->sp_key_smem != smem) {
    int per_sm = 0;
    if (dense_deep(smem)) {
      TRY(set_smem(h, k_num_dense_cta<T, true, 12>, smem));
      CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_num_dense_cta<T, true, 12>, nt, smem));
    } else {
      TRY(set_smem(h, k_num_dense_cta<T, true, 8>, smem));
      CUDA_TRY(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_num_dense_cta<T, true, 8>, nt, smem));
    }
    if (per_sm < 1) return fail(h, B200SP_ERR_CUDA, "k_num_dense_cta<single pass> does not fit an SM");
    h->sp_key_t = key_t; h->sp_key_nt = nt; h->sp_key_smem = smem; h->sp_per_sm = per_sm;
  }
  const int grid = std::max(1, std::min(p.m, h->num_sms * h->sp_per_sm));
  CUDA_TRY(h, cudaMemsetAsync(status, 0, (size_t)p.m * 8, st));
  CUDA_TRY(h, cudaMemsetAsync(&ctr->num_cursor[0], 0, sizeof(int), st));  // the row ticket
  SinglePassArgs sp{&ctr->num_cursor[0], status, p.Cp, &ctr->nnz_c, guarded ? &ctr->overflow : nullptr};
  const PBEnt<T> *pb = nullptr;
  int pb_last = 0;
  TRY(pack_b<T>(h, (const T *)Bx, &pb, &pb_last));
  if (dense_deep(smem))
    k_num_dense_cta<T, true, 12><<<grid, nt, smem, st>>>(nullptr, p.m, p.n, p.Ap, p.Aj, (const T *)Ax, p.Bp, pb, pb_last, nullptr, Cj,
                                                        (T *)Cx, a, sp);
  else
    k_num_dense_cta<T, true, 8><<<grid, nt, smem, st>>>(nullptr, p.m, p.n, p.Ap, p.Aj, (const T *)Ax, p.Bp, pb, pb_last, nullptr, Cj,
                                                       (T *)Cx, a, sp);
  return check_launch(h, "k_num_dense_cta<single pass>");
}

}  // namespace

extern "C" int b200sp_spgemm_single_pass(b200sp_handle_t h, int dtype, const void *alpha, const void *dAx,
                                         const void *dBx, int32_t *dCj, void *dCx, int64_t capacity, int64_t *nnzC) {
  if (!h) return B200SP_ERR_INVALID;
  Plan &p = h->plan;
  if (!p.stage1 || !p.single_pass)
    return fail(h, B200SP_ERR_STATE, "spgemm_single_pass needs a b200sp_spgemm_estimate that returned single_pass = 1");
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  if (!nnzC) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  if ((size_t)p.n * (dtype == B200SP_R_64F ? 8 : 4) > SMEM_DENSE_MAX)
    return fail(h, B200SP_ERR_INVALID, "dtype differs from the one given to b200sp_spgemm_estimate");
  if (capacity < p.nnz_bound)
    return fail(h, B200SP_ERR_INVALID, "capacity %lld is below the bound %lld of nnz(C)", (long long)capacity,
                (long long)p.nnz_bound);
  *nnzC = 0;
  DeviceGuard guard(h->device);
  cudaStream_t st = h->stream;
  p.stage1 = false;  // consumed
  if (p.nnz_bound == 0) {  // no products at all: C is empty
    CUDA_TRY(h, cudaMemsetAsync(p.Cp, 0, (size_t)(p.m + 1) * sizeof(int), st));
    cudaEventElapsedTime(&h->ms_symbolic, h->ev[0], h->ev[1]);
    h->ms_numeric = 0.f;
    p.nnzC = 0;
    p.valid = false;
    return B200SP_OK;
  }
  if (!dAx || !dBx || !dCj || !dCx) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  cudaEventRecord(h->ev[2], st);
  int rc = dtype == B200SP_R_32F ? run_single_pass<float>(h, alpha, dAx, dBx, dCj, dCx)
                                 : run_single_pass<double>(h, alpha, dAx, dBx, dCj, dCx);
  cudaEventRecord(h->ev[3], st);
  if (rc != B200SP_OK) return rc;
  // the one other host synchronisation: nnz(C) (the reference's cusparseSpMatGetSize sync)
  Counters *ctr = (Counters *)h->ws.ptr;
  CUDA_TRY(h, cudaMemcpyAsync(&h->h_counters->nnz_c, &ctr->nnz_c, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaStreamSynchronize(st));
  cudaEventElapsedTime(&h->ms_symbolic, h->ev[0], h->ev[1]);
  h->numeric_timed = true;
  p.nnzC = (int64_t)h->h_counters->nnz_c;
  *nnzC = p.nnzC;
  p.valid = false;  // no row lists exist: a value-only re-run has to go through estimate again
  return B200SP_OK;
}

extern "C" int b200sp_spgemm_single_pass_direct(b200sp_handle_t h, int dtype, int64_t m, int64_t k, int64_t n,
                                                int64_t nnzA, int64_t nnzB, const int32_t *dAp, const int32_t *dAj,
                                                const void *dAx, const int32_t *dBp, const int32_t *dBj,
                                                const void *dBx, const void *alpha, int32_t *dCp, int32_t *dCj,
                                                void *dCx, int64_t capacity, int64_t *nnzC, int64_t *num_products) {
  if (!h) return B200SP_ERR_INVALID;
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  if (!nnzC) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  *nnzC = 0;
  if (num_products) *num_products = 0;
  if (m < 0 || n < 0 || (size_t)n * (dtype == B200SP_R_64F ? 8 : 4) > SMEM_DENSE_MAX)
    return fail(h, B200SP_ERR_INVALID, "single_pass_direct: an n-column accumulator does not fit shared memory (n = %lld)", (long long)n);
  if (m * n > (int64_t)INT32_MAX || capacity < m * n)
    return fail(h, B200SP_ERR_INVALID, "single_pass_direct: capacity %lld is below m * n = %lld (or m * n exceeds int32)",
                (long long)capacity, (long long)(m * n));
  DeviceGuard guard(h->device);
  TRY(spgemm_stage1(h, m, k, n, nnzA, nnzB, dAp, dAj, dBp, dBj, B200SP_ALG_DEFAULT, 0.2f, dCp, true));
  Plan &p = h->plan;
  cudaStream_t st = h->stream;
  if (m == 0) return B200SP_OK;
  if (nnzA == 0 || nnzB == 0) {  // no products: C is empty
    CUDA_TRY(h, cudaMemsetAsync(dCp, 0, (size_t)(m + 1) * sizeof(int), st));
    CUDA_TRY(h, cudaStreamSynchronize(st));
    return B200SP_OK;
  }
  if (!dAx || !dBx || !dCj || !dCx) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  // C.indptr of rows without entries is written by the kernel too; rows are all visited
  cudaEventRecord(h->ev[2], st);
  int rc = dtype == B200SP_R_32F ? run_single_pass<float>(h, alpha, dAx, dBx, dCj, dCx, true)
                                 : run_single_pass<double>(h, alpha, dAx, dBx, dCj, dCx, true);
  cudaEventRecord(h->ev[3], st);
  if (rc != B200SP_OK) return rc;
  Counters *ctr = (Counters *)h->ws.ptr;
  CUDA_TRY(h, cudaMemcpyAsync(h->h_counters, ctr, sizeof(Counters), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaStreamSynchronize(st));
  cudaEventElapsedTime(&h->ms_symbolic, h->ev[0], h->ev[1]);
  h->numeric_timed = true;
  if (h->h_counters->overflow)
    return fail(h, B200SP_ERR_OVERFLOW, "a row of A@B has more than 2^31-1 intermediate products");
  if (h->h_counters->b_bad)
    return fail(h, B200SP_ERR_INVALID,
                "B is not canonical (a row has unsorted, duplicate or out-of-range column indices): "
                "canonicalise it first (b200sp_csr_sum_duplicates / b200sp_csr_sort_indices)");
  p.num_products = (int64_t)h->h_counters->total_products;
  p.max_row_products = h->h_counters->max_row_products;
  p.nnz_bound = (int64_t)h->h_counters->total_ub;
  p.nnzC = (int64_t)h->h_counters->nnz_c;
  if (num_products) *num_products = p.num_products;
  *nnzC = p.nnzC;
  p.valid = false;
  return B200SP_OK;
}

extern "C" int b200sp_spgemm_numeric(b200sp_handle_t h, int dtype, const void *alpha, const void *dAx,
                                     const void *dBx, int32_t *dCj, void *dCx) {
  if (!h) return B200SP_ERR_INVALID;
  Plan &p = h->plan;
  if (!p.valid) return fail(h, B200SP_ERR_STATE, "spgemm_numeric called without a successful spgemm_symbolic");
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  if (p.nnzC == 0) return B200SP_OK;
  if (!dAx || !dBx || !dCj || !dCx) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  DeviceGuard guard(h->device);
  cudaEventRecord(h->ev[2], h->stream);
  int rc = dtype == B200SP_R_32F ? run_numeric<float>(h, alpha, dAx, dBx, dCj, dCx)
                                 : run_numeric<double>(h, alpha, dAx, dBx, dCj, dCx);
  cudaEventRecord(h->ev[3], h->stream);
  h->numeric_timed = (rc == B200SP_OK);
  return rc;
}

// ========================================================================== SpMV / SpMM ====

namespace {

template <typename T>
int run_spmv(b200sp_context *h, int m, int n, int64_t nnz, int64_t ncols, const int *Ap, const int *Aj,
             const T *Ax, const T *X, int64_t ldx, T *Y, int64_t ldy, const void *alpha, const void *beta) {
  (void)n;
  const T a = alpha ? *(const T *)alpha : T(1);
  const T b = beta ? *(const T *)beta : T(0);
  cudaStream_t st = h->stream;
  if (m == 0 || ncols == 0) return B200SP_OK;
  // Kernel choice.  Row-group ("vector") kernel with 128-bit loads for matrices whose rows are of
  // similar length (C2: 67 us vs 75 us merge-path); merge-path when one row can hold a warp up
  // (R-MAT scale 20, longest row 39 705: 89 us vs 1875 us).  The longest row is measured once per
  // matrix (one small kernel + a 4-byte D2H) and cached on the handle.
  // B200SP_SPMV_KERNEL = vector | vector4 | merge forces one (vector4 = the 128-bit row-group kernel
  // whatever the row length); "+scalar" (e.g. "merge+scalar") forces the 4-byte-load variants (what
  // bench.py uses to report all of them).
  const char *mode = getenv("B200SP_SPMV_KERNEL");
  const bool force_vector = mode && strncmp(mode, "vector", 6) == 0;
  const bool force_merge = mode && (strncmp(mode, "merge", 5) == 0 || strncmp(mode, "tile", 4) == 0);
  const bool scalar_only = mode && strstr(mode, "scalar") != nullptr;
  const bool force_vec4 = mode && strncmp(mode, "vector4", 7) == 0;
  const bool aligned = (((uintptr_t)Aj | (uintptr_t)Ax) & 15) == 0;
  const bool vec = aligned && !scalar_only;
  bool use_merge = ncols == 1 && nnz > 0;
  if (use_merge && force_vector) use_merge = false;
  if (use_merge && !force_merge) {
    if (h->spmv_key_ap != Ap || h->spmv_key_m != m || h->spmv_key_nnz != nnz) {
      TRY(ensure(h, h->ws2, 256));
      int *d_max = (int *)h->ws2.ptr;
      CUDA_TRY(h, cudaMemsetAsync(d_max, 0, sizeof(int), st));
      k_max_row_len<<<std::max(1, std::min((m + 255) / 256, h->num_sms * 8)), 256, 0, st>>>(m, Ap, d_max);
      TRY(check_launch(h, "k_max_row_len"));
      int mx = 0;
      CUDA_TRY(h, cudaMemcpyAsync(&mx, d_max, sizeof(int), cudaMemcpyDeviceToHost, st));
      CUDA_TRY(h, cudaStreamSynchronize(st));
      h->spmv_key_ap = Ap;
      h->spmv_key_m = m;
      h->spmv_key_nnz = nnz;
      h->spmv_max_row = mx;
    }
    const double avg_len = (double)nnz / (double)m;
    // a row-group kernel finishes when its longest row does: fine while that row is a few trips
    use_merge = (double)h->spmv_max_row > 32.0 * avg_len + 512.0;
  }
  if (use_merge) {
    const long long total = (long long)m + nnz;  // merge items: row-end markers + nonzeros
    const int ntiles = (int)((total + SPMV_TILE - 1) / SPMV_TILE);
    const size_t o_i = align_up((size_t)(ntiles + 1) * sizeof(int), 256);
    const size_t o_h = o_i + align_up((size_t)ntiles * sizeof(SpmvTileInfo), 256);
    const size_t o_c = o_h + align_up((size_t)ntiles * sizeof(T), 256);
    TRY(ensure(h, h->ws2, o_c + (size_t)ntiles * sizeof(T)));
    char *base = (char *)h->ws2.ptr;
    int *tile_row = (int *)base;
    SpmvTileInfo *info = (SpmvTileInfo *)(base + o_i);
    T *head = (T *)(base + o_h), *carry = (T *)(base + o_c);
    k_spmv_partition<<<(ntiles + 1 + 7) / 8, 256, 0, st>>>(m, total, ntiles, Ap, tile_row);
    TRY(check_launch(h, "k_spmv_partition"));
    if (vec) k_spmv_merge<T, true><<<ntiles, SPMV_THREADS, 0, st>>>(m, (int)nnz, tile_row, Ap, Aj, Ax, X, Y, a, b, info, head, carry);
    else k_spmv_merge<T, false><<<ntiles, SPMV_THREADS, 0, st>>>(m, (int)nnz, tile_row, Ap, Aj, Ax, X, Y, a, b, info, head, carry);
    TRY(check_launch(h, "k_spmv_merge"));
    if (ntiles == 1) return B200SP_OK;  // no row can span tiles
    k_spmv_fixup<T><<<(ntiles + 255) / 256, 256, 0, st>>>(ntiles, info, head, carry, Y, a, b);
    return check_launch(h, "k_spmv_fixup");
  }
  const double avg = (double)nnz / (double)m;
  // 128-bit loads pay only once a row fills all 32 lanes several times over: measured on B200 with
  // 1.1e7 nonzeros (profiles/r02_spmv_sweep_v16.jsonl, us, 4-byte / 128-bit loads): 10.5 entries per row
  // 67.6 / 74.6, 64: 60.3 / 66.4, 128: 51.2 / 55.3, 256: 45.1 / 38.9 (unaligned row starts mask half of
  // every 16-byte load of a short row)
  if (vec && (force_vec4 || avg >= 192.0)) {  // four entries per lane and trip: a quarter of the lanes per row
    const int G = avg <= 3 ? 1 : avg <= 6 ? 2 : avg <= 12 ? 4 : avg <= 24 ? 8 : avg <= 48 ? 16 : 32;
    const int gx = std::max(1, std::min((int)(((int64_t)m + 256 / G - 1) / (256 / G)), h->num_sms * 16));
    dim3 grid(gx, (unsigned)ncols);
    switch (G) {
      case 1: k_spmv_vector4<1, T><<<grid, 256, 0, st>>>(m, (int)nnz, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
      case 2: k_spmv_vector4<2, T><<<grid, 256, 0, st>>>(m, (int)nnz, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
      case 4: k_spmv_vector4<4, T><<<grid, 256, 0, st>>>(m, (int)nnz, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
      case 8: k_spmv_vector4<8, T><<<grid, 256, 0, st>>>(m, (int)nnz, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
      case 16: k_spmv_vector4<16, T><<<grid, 256, 0, st>>>(m, (int)nnz, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
      default: k_spmv_vector4<32, T><<<grid, 256, 0, st>>>(m, (int)nnz, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
    }
    return check_launch(h, "k_spmv_vector4");
  }
  const int G = avg <= 3 ? 2 : avg <= 6 ? 4 : avg <= 12 ? 8 : avg <= 24 ? 16 : 32;
  const int gx = std::max(1, std::min((int)(((int64_t)m + 256 / G - 1) / (256 / G)), h->num_sms * 16));
  dim3 grid(gx, (unsigned)ncols);
  switch (G) {
    case 2: k_spmv_vector<2, T><<<grid, 256, 0, st>>>(m, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
    case 4: k_spmv_vector<4, T><<<grid, 256, 0, st>>>(m, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
    case 8: k_spmv_vector<8, T><<<grid, 256, 0, st>>>(m, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
    case 16: k_spmv_vector<16, T><<<grid, 256, 0, st>>>(m, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
    default: k_spmv_vector<32, T><<<grid, 256, 0, st>>>(m, Ap, Aj, Ax, X, ldx, Y, ldy, a, b); break;
  }
  return check_launch(h, "k_spmv_vector");
}

}  // namespace

extern "C" int b200sp_spmm(b200sp_handle_t h, int dtype, int64_t m, int64_t n, int64_t nnz, int64_t ncols,
                           const int32_t *dAp, const int32_t *dAj, const void *dAx, const void *dB,
                           int64_t ldb, void *dC, int64_t ldc, const void *alpha, const void *beta) {
  if (!h) return B200SP_ERR_INVALID;
  if (m < 0 || n < 0 || nnz < 0 || ncols < 0 || m >= INT32_MAX || n >= INT32_MAX || nnz > INT32_MAX || ncols > 65535)
    return fail(h, B200SP_ERR_INVALID, "bad shape m=%lld n=%lld nnz=%lld ncols=%lld", (long long)m, (long long)n,
                (long long)nnz, (long long)ncols);
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  if (m == 0 || ncols == 0) return B200SP_OK;
  if (!dAp || !dC || (nnz && (!dAj || !dAx || !dB))) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  if (ncols > 1 && (ldb < n || ldc < m)) return fail(h, B200SP_ERR_INVALID, "leading dimension too small");
  DeviceGuard guard(h->device);
  if (dtype == B200SP_R_32F)
    return run_spmv<float>(h, (int)m, (int)n, nnz, ncols, dAp, dAj, (const float *)dAx, (const float *)dB, ldb,
                           (float *)dC, ldc, alpha, beta);
  return run_spmv<double>(h, (int)m, (int)n, nnz, ncols, dAp, dAj, (const double *)dAx, (const double *)dB, ldb,
                          (double *)dC, ldc, alpha, beta);
}

extern "C" int b200sp_spmv(b200sp_handle_t h, int dtype, int64_t m, int64_t n, int64_t nnz, const int32_t *dAp,
                           const int32_t *dAj, const void *dAx, const void *dx, void *dy, const void *alpha,
                           const void *beta) {
  return b200sp_spmm(h, dtype, m, n, nnz, 1, dAp, dAj, dAx, dx, n, dy, m, alpha, beta);
}

// ----------------------------------------------------- SpMV without any allocation ----
// Merge-path SpMV whose per-tile scratch (tile boundary rows, tile info, head / carry partials)
// lives in a file-scope device array instead of the handle's workspace, for callers that must not
// reach cudaMalloc (stream capture, serving loops).  Every GPU that loads the library holds the
// array (B200SP_NOALLOC_SCRATCH_BYTES) until exit.  All calls on a device share it: calls that
// overlap on the GPU (two streams, two handles, two threads) race.  Nothing in it has to start at
// zero - k_spmv_partition writes every boundary, k_spmv_merge writes info[k] of every tile, and
// k_spmv_fixup reads head / carry entries only where this launch's info says they were written -
// so the previous call's contents are simply overwritten.
namespace {

constexpr size_t NOALLOC_O_INFO = align_up_c((size_t)(B200SP_NOALLOC_MAX_TILES + 1) * sizeof(int), 256);
constexpr size_t NOALLOC_O_HEAD = NOALLOC_O_INFO + align_up_c((size_t)B200SP_NOALLOC_MAX_TILES * sizeof(b200sp::SpmvTileInfo), 256);
constexpr size_t NOALLOC_O_CARRY = NOALLOC_O_HEAD + align_up_c((size_t)B200SP_NOALLOC_MAX_TILES * sizeof(double), 256);
constexpr size_t NOALLOC_BYTES = NOALLOC_O_CARRY + align_up_c((size_t)B200SP_NOALLOC_MAX_TILES * sizeof(double), 256);
static_assert(NOALLOC_BYTES == B200SP_NOALLOC_SCRATCH_BYTES, "include/b200sp.h states the size of the scratch array");

}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

namespace {

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
void *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};  // device address of g_noalloc_scratch, per device

// In host code `g_noalloc_scratch` is only the host-side shadow of the symbol; the address a
// kernel may use comes from cudaGetSymbolAddress with the owning device current (resolved once per
// device, at the first call).
int noalloc_scratch(b200sp_context *h, char **out) {
  if (h->device < 0 || h->device >= NOALLOC_MAX_DEVICES) return fail(h, B200SP_ERR_INVALID, "device %d out of range", h->device);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[h->device]) {
    int cur = -1;
    CUDA_TRY(h, cudaGetDevice(&cur));
    if (cur != h->device) return fail(h, B200SP_ERR_STATE, "device %d is not current (%d is)", h->device, cur);
    void *p = nullptr;
    CUDA_TRY(h, cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) return fail(h, B200SP_ERR_CUDA, "cudaGetSymbolAddress returned NULL");
    g_noalloc_ptr[h->device] = p;
  }
  *out = (char *)g_noalloc_ptr[h->device];
  return B200SP_OK;
}

template <typename T>
int run_spmv_noalloc(b200sp_context *h, int m, int64_t nnz, const int *Ap, const int *Aj, const T *Ax, const T *x,
                     T *y, const void *alpha, const void *beta) {
  const T a = alpha ? *(const T *)alpha : T(1);
  const T b = beta ? *(const T *)beta : T(0);
  const long long total = (long long)m + nnz;
  const long long ntiles_ll = (total + SPMV_TILE - 1) / SPMV_TILE;
  if (ntiles_ll > B200SP_NOALLOC_MAX_TILES)
    return fail(h, B200SP_ERR_INVALID,
                "b200sp_spmv_noalloc: m + nnz = %lld exceeds the fixed scratch (%lld merge items = %d tiles of %d); "
                "use b200sp_spmv", total, (long long)B200SP_NOALLOC_MAX_TILES * SPMV_TILE, B200SP_NOALLOC_MAX_TILES,
                SPMV_TILE);
  const int ntiles = (int)ntiles_ll;
  char *base = nullptr;
  TRY(noalloc_scratch(h, &base));
  int *tile_row = (int *)base;
  SpmvTileInfo *info = (SpmvTileInfo *)(base + NOALLOC_O_INFO);
  T *head = (T *)(base + NOALLOC_O_HEAD), *carry = (T *)(base + NOALLOC_O_CARRY);
  cudaStream_t st = h->stream;
  const bool vec = (((uintptr_t)Aj | (uintptr_t)Ax) & 15) == 0;
  k_spmv_partition<<<(ntiles + 1 + 7) / 8, 256, 0, st>>>(m, total, ntiles, Ap, tile_row);
  TRY(check_launch(h, "k_spmv_partition"));
  if (vec) k_spmv_merge<T, true><<<ntiles, SPMV_THREADS, 0, st>>>(m, (int)nnz, tile_row, Ap, Aj, Ax, x, y, a, b, info, head, carry);
  else k_spmv_merge<T, false><<<ntiles, SPMV_THREADS, 0, st>>>(m, (int)nnz, tile_row, Ap, Aj, Ax, x, y, a, b, info, head, carry);
  TRY(check_launch(h, "k_spmv_merge"));
  if (ntiles == 1) return B200SP_OK;
  k_spmv_fixup<T><<<(ntiles + 255) / 256, 256, 0, st>>>(ntiles, info, head, carry, y, a, b);
  return check_launch(h, "k_spmv_fixup");
}

}  // namespace

extern "C" int b200sp_spmv_noalloc(b200sp_handle_t h, int dtype, int64_t m, int64_t n, int64_t nnz,
                                   const int32_t *dAp, const int32_t *dAj, const void *dAx, const void *dx,
                                   void *dy, const void *alpha, const void *beta) {
  if (!h) return B200SP_ERR_INVALID;
  if (m < 0 || n < 0 || nnz < 0 || m >= INT32_MAX || n >= INT32_MAX || nnz > INT32_MAX)
    return fail(h, B200SP_ERR_INVALID, "bad shape m=%lld n=%lld nnz=%lld", (long long)m, (long long)n, (long long)nnz);
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  if (m == 0) return B200SP_OK;
  if (!dAp || !dy || (nnz && (!dAj || !dAx || !dx))) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  DeviceGuard guard(h->device);
  if (dtype == B200SP_R_32F)
    return run_spmv_noalloc<float>(h, (int)m, nnz, dAp, dAj, (const float *)dAx, (const float *)dx, (float *)dy, alpha, beta);
  return run_spmv_noalloc<double>(h, (int)m, nnz, dAp, dAj, (const double *)dAx, (const double *)dx, (double *)dy, alpha, beta);
}

// ==================================================================== canonical helpers ====

extern "C" int b200sp_csr_is_canonical(b200sp_handle_t h, int64_t m, int64_t ncols, const int32_t *dAp,
                                       const int32_t *dAj, int *result) {
  if (!h || !result || m < 0 || m >= INT32_MAX || ncols < 0 || ncols > INT32_MAX || !dAp)
    return h ? fail(h, B200SP_ERR_INVALID, "bad argument") : B200SP_ERR_INVALID;
  *result = 1;
  if (m == 0) return B200SP_OK;
  DeviceGuard guard(h->device);
  TRY(ensure(h, h->ws2, 256));
  int *flag = (int *)h->ws2.ptr;
  CUDA_TRY(h, cudaMemsetAsync(flag, 0, sizeof(int), h->stream));
  const int grid = std::max(1, std::min((int)((m + 7) / 8), h->num_sms * 8));
  k_csr_check_canonical<<<grid, 256, 0, h->stream>>>((int)m, (int)ncols, dAp, dAj, flag);
  TRY(check_launch(h, "k_csr_check_canonical"));
  int bad = 0;
  CUDA_TRY(h, cudaMemcpyAsync(&bad, flag, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(h, cudaStreamSynchronize(h->stream));
  *result = bad ? 0 : 1;
  return B200SP_OK;
}

namespace {
template <typename T>
int run_sort_indices_stable(b200sp_context *h, int64_t m, int64_t nnz, const int *Ap, int *Aj, T *Ax);  // convert_api.inl
}

extern "C" int b200sp_csr_sort_indices(b200sp_handle_t h, int dtype, int64_t m, int64_t nnz, const int32_t *dAp,
                                       int32_t *dAj, void *dAx) {
  if (!h) return B200SP_ERR_INVALID;
  if (m < 0 || m >= INT32_MAX || nnz < 0 || nnz > INT32_MAX || !dAp) return fail(h, B200SP_ERR_INVALID, "bad argument");
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  if (m == 0 || nnz == 0) return B200SP_OK;
  DeviceGuard guard(h->device);
  TRY(ensure(h, h->ws2, 256));
  int *flag = (int *)h->ws2.ptr;
  CUDA_TRY(h, cudaMemsetAsync(flag, 0, sizeof(int), h->stream));
  const size_t vs = dtype == B200SP_R_32F ? 4 : 8;
  const int cap = 16384;
  const size_t smem = (size_t)cap * (vs + 4);
  const int grid = std::max(1, std::min((int)m, h->num_sms * 2));
  if (dtype == B200SP_R_32F) {
    TRY(set_smem(h, k_csr_sort_rows<float>, smem));
    k_csr_sort_rows<float><<<grid, 256, smem, h->stream>>>((int)m, cap, dAp, dAj, (float *)dAx, flag);
  } else {
    TRY(set_smem(h, k_csr_sort_rows<double>, smem));
    k_csr_sort_rows<double><<<grid, 256, smem, h->stream>>>((int)m, cap, dAp, dAj, (double *)dAx, flag);
  }
  TRY(check_launch(h, "k_csr_sort_rows"));
  int too_long = 0;
  CUDA_TRY(h, cudaMemcpyAsync(&too_long, flag, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(h, cudaStreamSynchronize(h->stream));
  if (!too_long) return B200SP_OK;
  // a row longer than the shared-memory sort, or duplicate columns (whose order must be kept): stable radix sort
  return dtype == B200SP_R_32F ? run_sort_indices_stable<float>(h, m, nnz, dAp, dAj, (float *)dAx)
                               : run_sort_indices_stable<double>(h, m, nnz, dAp, dAj, (double *)dAx);
}

// ===================================== canonicalisation, dense, compare ====

#include "convert_api.inl"

// ================================================================= host-buffer one-shots ====

namespace {

int ensure_pinned(b200sp_context *h, int i, size_t bytes) {
  if (bytes <= h->pinned_out_size[i]) return B200SP_OK;
  if (h->pinned_out[i]) cudaFreeHost(h->pinned_out[i]);
  h->pinned_out[i] = nullptr;
  h->pinned_out_size[i] = 0;
  const size_t want = std::max<size_t>(align_up(bytes, 4096), 4096);
  if (cudaMallocHost(&h->pinned_out[i], want) != cudaSuccess) {
    cudaGetLastError();
    return fail(h, B200SP_ERR_ALLOC, "out of pinned host memory allocating %zu bytes", want);
  }
  h->pinned_out_size[i] = want;
  return B200SP_OK;
}

int upload_on(b200sp_context *h, cudaStream_t st, Buffer &b, const void *src, size_t bytes) {
  TRY(ensure(h, b, std::max<size_t>(bytes, 16)));
  if (bytes) CUDA_TRY(h, cudaMemcpyAsync(b.ptr, src, bytes, cudaMemcpyHostToDevice, st));
  return B200SP_OK;
}

int upload(b200sp_context *h, Buffer &b, const void *src, size_t bytes) { return upload_on(h, h->stream, b, src, bytes); }

}  // namespace

extern "C" int b200sp_spgemm_host(b200sp_handle_t h, int dtype, int64_t m, int64_t k, int64_t n,
                                  const int32_t *hAp, const int32_t *hAj, const void *hAx,
                                  const int32_t *hBp, const int32_t *hBj, const void *hBx, int alg,
                                  float chunk_fraction, const void *alpha, int32_t **hCp, int32_t **hCj,
                                  void **hCx, int64_t *nnzC, int64_t *num_products) {
  if (!h) return B200SP_ERR_INVALID;
  if (!hAp || !hBp || !hCp || !hCj || !hCx || !nnzC || m < 0 || k < 0 || n < 0)
    return fail(h, B200SP_ERR_INVALID, "bad argument");
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  const size_t vs = dtype == B200SP_R_32F ? 4 : 8;
  const int64_t nnzA = hAp[m], nnzB = hBp[k];
  if (nnzA < 0 || nnzB < 0 || (nnzA && (!hAj || !hAx)) || (nnzB && (!hBj || !hBx)))
    return fail(h, B200SP_ERR_INVALID, "bad CSR arrays");
  DeviceGuard guard(h->device);
  const bool alias = (hAp == hBp && hAj == hBj && hAx == hBx && m == k);
  // Ingest pipeline (the reference times H2D inside the op:
  // SpGEMM_alg_comparison/profiler.py:210-213).  The symbolic pass needs only the structure
  // arrays, so those go first on the compute stream while the value arrays travel on a second
  // DMA queue underneath the symbolic kernels; the numeric pass waits for them with an event.
  // C.indptr is final after the symbolic pass and goes back under the numeric kernels.
  cudaStream_t cs = h->copy_stream ? h->copy_stream : h->stream;
  TRY(upload(h, h->io[0], hAp, (size_t)(m + 1) * 4));
  TRY(upload(h, h->io[1], hAj, (size_t)nnzA * 4));
  if (!alias) {
    TRY(upload(h, h->io[3], hBp, (size_t)(k + 1) * 4));
    TRY(upload(h, h->io[4], hBj, (size_t)nnzB * 4));
  }
  TRY(upload_on(h, cs, h->io[2], hAx, (size_t)nnzA * vs));
  if (!alias) TRY(upload_on(h, cs, h->io[5], hBx, (size_t)nnzB * vs));
  CUDA_TRY(h, cudaEventRecord(h->ev_copy, cs));
  const int bi = alias ? 0 : 3;
  TRY(ensure(h, h->io[6], (size_t)(m + 1) * 4));
  int64_t nc = 0;
  int rc = b200sp_spgemm_symbolic(h, m, k, n, nnzA, nnzB, (const int *)h->io[0].ptr, (const int *)h->io[1].ptr,
                                  (const int *)h->io[bi].ptr, (const int *)h->io[bi + 1].ptr, alg, chunk_fraction,
                                  num_products, &nc, (int *)h->io[6].ptr);
  if (rc != B200SP_OK) {
    cudaStreamSynchronize(cs);  // the caller may free its value arrays as soon as we return
    return rc;
  }
  // ALG3 = the memory-bounded mode (spgemm_from_txt_alg3.cu:101-115, 195-208: chunk_fraction sizes
  // the compute buffer).  ALG1 already holds nothing but inputs + C + 12 m bytes, so what is left
  // to bound is C itself: the rows are cut into chunks of at most chunk_fraction * nnz(C) entries,
  // chunk c + 1 is computed while chunk c travels to the host, and the device only ever holds two
  // chunk-sized output buffers.  The price is a second structure pass per chunk (the plan of the
  // chunk's rows) and 2 host synchronisations per chunk.  ALG2 runs the ALG1 path.
  const bool chunked = alg == B200SP_ALG3 && nc > 0 && chunk_fraction < 1.f && m > 1;
  rc = ensure_pinned(h, 0, (size_t)(m + 1) * 4);
  if (rc == B200SP_OK) rc = ensure_pinned(h, 1, (size_t)nc * 4);
  if (rc == B200SP_OK) rc = ensure_pinned(h, 2, (size_t)nc * vs);
  if (rc == B200SP_OK && !chunked) rc = ensure(h, h->io[7], std::max<size_t>((size_t)nc * 4, 16));
  if (rc == B200SP_OK && !chunked) rc = ensure(h, h->io[8], std::max<size_t>((size_t)nc * vs, 16));
  if (rc != B200SP_OK) {
    cudaStreamSynchronize(cs);
    return rc;
  }
  // C.indptr is final once the symbolic pass is done
  CUDA_TRY(h, cudaEventRecord(h->ev_sym, h->stream));
  CUDA_TRY(h, cudaStreamWaitEvent(cs, h->ev_sym, 0));
  CUDA_TRY(h, cudaMemcpyAsync(h->pinned_out[0], h->io[6].ptr, (size_t)(m + 1) * 4, cudaMemcpyDeviceToHost, cs));
  CUDA_TRY(h, cudaStreamWaitEvent(h->stream, h->ev_copy, 0));
  if (chunked) {
    CUDA_TRY(h, cudaStreamSynchronize(cs));  // the host cuts the chunks from C.indptr
    const int32_t *cp = (const int32_t *)h->pinned_out[0];
    const int64_t total_products = h->plan.num_products;
    const int64_t budget = std::max<int64_t>(1, (int64_t)std::floor((double)chunk_fraction * (double)nc));
    std::vector<int64_t> cuts(1, 0);
    int64_t max_chunk = 0, max_rows = 0;
    for (int64_t r0 = 0; r0 < m;) {
      // last r1 with cp[r1] - cp[r0] <= budget; a row larger than the budget is a chunk of its own
      int64_t r1 = std::upper_bound(cp + r0 + 1, cp + m + 1, (int32_t)std::min<int64_t>(cp[r0] + budget, INT32_MAX)) - cp - 1;
      r1 = std::min<int64_t>(std::max<int64_t>(r1, r0 + 1), m);
      cuts.push_back(r1);
      max_chunk = std::max<int64_t>(max_chunk, cp[r1] - cp[r0]);
      max_rows = std::max<int64_t>(max_rows, r1 - r0);
      r0 = r1;
    }
    rc = ensure(h, h->io[7], std::max<size_t>((size_t)max_chunk * 4 * 2, 16));
    if (rc == B200SP_OK) rc = ensure(h, h->io[8], std::max<size_t>((size_t)max_chunk * vs * 2, 16));
    if (rc == B200SP_OK) rc = ensure(h, h->io[9], (size_t)(max_rows + 1) * 4);
    if (rc != B200SP_OK) return rc;
    const int *dAp = (const int *)h->io[0].ptr;
    for (size_t c = 0; c + 1 < cuts.size(); ++c) {
      const int64_t r0 = cuts[c], r1 = cuts[c + 1];
      const int64_t want = (int64_t)cp[r1] - cp[r0];
      if (want == 0) continue;
      const int slot = (int)(c & 1);
      int32_t *dj = (int32_t *)h->io[7].ptr + (size_t)slot * max_chunk;
      char *dx = (char *)h->io[8].ptr + (size_t)slot * max_chunk * vs;
      int64_t got = 0, np_chunk = 0;
      // the row panel [r0, r1) of A: same index / value arrays, shifted row pointers (the kernels
      // only ever read Ap[row] and Ap[row + 1] as absolute positions)
      rc = b200sp_spgemm_symbolic(h, r1 - r0, k, n, (int64_t)hAp[r1] - hAp[r0], nnzB, dAp + r0, (const int *)h->io[1].ptr,
                                  (const int *)h->io[bi].ptr, (const int *)h->io[bi + 1].ptr, B200SP_ALG1, 1.f,
                                  &np_chunk, &got, (int *)h->io[9].ptr);
      if (rc == B200SP_OK && got != want)
        rc = fail(h, B200SP_ERR_STATE, "chunk [%lld,%lld): structure pass gave %lld entries, C.indptr says %lld",
                  (long long)r0, (long long)r1, (long long)got, (long long)want);
      if (rc != B200SP_OK) {
        cudaStreamSynchronize(cs);
        return rc;
      }
      CUDA_TRY(h, cudaStreamWaitEvent(h->stream, h->ev_slot[slot], 0));  // chunk c - 2 has left this buffer
      rc = b200sp_spgemm_numeric(h, dtype, alpha, h->io[2].ptr, h->io[bi + 2].ptr, dj, dx);
      if (rc != B200SP_OK) {
        cudaStreamSynchronize(cs);
        return rc;
      }
      CUDA_TRY(h, cudaEventRecord(h->ev_chunk, h->stream));
      CUDA_TRY(h, cudaStreamWaitEvent(cs, h->ev_chunk, 0));
      CUDA_TRY(h, cudaMemcpyAsync((int32_t *)h->pinned_out[1] + cp[r0], dj, (size_t)want * 4, cudaMemcpyDeviceToHost, cs));
      CUDA_TRY(h, cudaMemcpyAsync((char *)h->pinned_out[2] + (size_t)cp[r0] * vs, dx, (size_t)want * vs, cudaMemcpyDeviceToHost, cs));
      CUDA_TRY(h, cudaEventRecord(h->ev_slot[slot], cs));
    }
    CUDA_TRY(h, cudaStreamSynchronize(cs));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    h->plan.num_products = total_products;  // report the whole product, not the last chunk
    h->plan.nnzC = nc;
    h->plan.num_chunks = (int)cuts.size() - 1;
    h->plan.valid = false;  // the plan left behind belongs to the last chunk only
    *hCp = (int32_t *)h->pinned_out[0];
    *hCj = (int32_t *)h->pinned_out[1];
    *hCx = h->pinned_out[2];
    *nnzC = nc;
    if (num_products) *num_products = total_products;
    return B200SP_OK;
  }
  rc = b200sp_spgemm_numeric(h, dtype, alpha, h->io[2].ptr, h->io[bi + 2].ptr, (int *)h->io[7].ptr, h->io[8].ptr);
  if (rc != B200SP_OK) {
    cudaStreamSynchronize(cs);
    return rc;
  }
  if (nc) {
    CUDA_TRY(h, cudaMemcpyAsync(h->pinned_out[1], h->io[7].ptr, (size_t)nc * 4, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(h->pinned_out[2], h->io[8].ptr, (size_t)nc * vs, cudaMemcpyDeviceToHost, h->stream));
  }
  CUDA_TRY(h, cudaStreamSynchronize(cs));
  CUDA_TRY(h, cudaStreamSynchronize(h->stream));
  *hCp = (int32_t *)h->pinned_out[0];
  *hCj = (int32_t *)h->pinned_out[1];
  *hCx = h->pinned_out[2];
  *nnzC = nc;
  return B200SP_OK;
}

extern "C" int b200sp_spmv_host(b200sp_handle_t h, int dtype, int64_t m, int64_t n, int64_t nnz,
                                const int32_t *hAp, const int32_t *hAj, const void *hAx, const void *hx,
                                void *hy, const void *alpha, const void *beta) {
  if (!h) return B200SP_ERR_INVALID;
  if (!hAp || !hy || m < 0 || n < 0 || nnz < 0 || (nnz && (!hAj || !hAx || !hx))) return fail(h, B200SP_ERR_INVALID, "bad argument");
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  const size_t vs = dtype == B200SP_R_32F ? 4 : 8;
  DeviceGuard guard(h->device);
  TRY(upload(h, h->io[0], hAp, (size_t)(m + 1) * 4));
  TRY(upload(h, h->io[1], hAj, (size_t)nnz * 4));
  TRY(upload(h, h->io[2], hAx, (size_t)nnz * vs));
  TRY(upload(h, h->io[3], hx, (size_t)n * vs));
  const bool beta_zero = !beta || (dtype == B200SP_R_32F ? *(const float *)beta == 0.f : *(const double *)beta == 0.0);
  if (beta_zero) TRY(ensure(h, h->io[4], std::max<size_t>((size_t)m * vs, 16)));
  else TRY(upload(h, h->io[4], hy, (size_t)m * vs));
  TRY(b200sp_spmv(h, dtype, m, n, nnz, (const int *)h->io[0].ptr, (const int *)h->io[1].ptr, h->io[2].ptr,
                  h->io[3].ptr, h->io[4].ptr, alpha, beta));
  if (m) CUDA_TRY(h, cudaMemcpyAsync(hy, h->io[4].ptr, (size_t)m * vs, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(h, cudaStreamSynchronize(h->stream));
  return B200SP_OK;
}

extern "C" int b200sp_free_host(void *ptr) {
  free(ptr);
  return B200SP_OK;
}

// ---- plumbing of the row-panel driver (spmm_b200/multi.py): one shared host buffer for C ----

extern "C" int b200sp_host_register(void *ptr, size_t bytes) {
  if (!ptr || !bytes) return B200SP_ERR_INVALID;
  const cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
  if (e == cudaErrorHostMemoryAlreadyRegistered) {
    cudaGetLastError();
    return B200SP_OK;
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? B200SP_ERR_ALLOC : B200SP_ERR_CUDA;
  }
  return B200SP_OK;
}

extern "C" int b200sp_host_unregister(void *ptr) {
  if (!ptr) return B200SP_ERR_INVALID;
  if (cudaHostUnregister(ptr) != cudaSuccess) {
    cudaGetLastError();
    return B200SP_ERR_CUDA;
  }
  return B200SP_OK;
}

extern "C" int b200sp_memcpy_async(b200sp_handle_t h, void *dst, const void *src, size_t bytes, int kind,
                                   void *cuda_stream) {
  if (!h || (kind != 0 && kind != 1)) return h ? fail(h, B200SP_ERR_INVALID, "bad argument") : B200SP_ERR_INVALID;
  if (!bytes) return B200SP_OK;
  if (!dst || !src) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  DeviceGuard guard(h->device);
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : (h->copy_stream ? h->copy_stream : h->stream);
  CUDA_TRY(h, cudaMemcpyAsync(dst, src, bytes, kind == 0 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, st));
  return B200SP_OK;
}
