/*
 * b200sp.h -- C ABI of libb200sp.so, the B200-native (sm_100a) CSR SpGEMM / SpMV engine.
 *
 * This header is the drop-in boundary for the one hot path BASELINE.json:north_star
 * names.  Every entry point states which call of the reference (wjustin784/spMM,
 * paths relative to the reference root) it takes the place of.  Plain pointers and
 * sizes only: no torch / CuPy types cross this boundary.
 *
 * Conventions
 *   - every function returns an int status: 0 = OK, negative = error; the text of
 *     the last error of a handle is b200sp_last_error(handle).  The library never
 *     calls exit() (the reference drivers do: cupy_cusparse/spgemm_from_txt_alg1.cu:12-17).
 *   - CSR is base-0, int32 indptr/indices, fp32 or fp64 values, rows sorted by column
 *     and duplicate free on input ("canonical", the precondition asserted at
 *     modify_src/cupy-src/cupyx/cusparse.py:2030-2031).  Output C is canonical too.
 *   - pointers named d* are DEVICE pointers on the handle's device, h* are HOST pointers.
 *   - A and B may alias (every reference profiler computes A @ A).
 *   - one handle per host thread / device; no global mutable state.
 *   - all work is enqueued on the handle's stream (default: the legacy default
 *     stream, like the reference: SpGEMM_alg_comparison/profiler.py:104-105).
 */
#ifndef B200SP_H_
#define B200SP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SP_VERSION 200 /* 0.2.0: round 2 (single-pass SpGEMM, packed B, bucket conversions, noalloc SpMV) */

typedef struct b200sp_context *b200sp_handle_t;

enum b200sp_status {
  B200SP_OK = 0,
  B200SP_ERR_INVALID = -1,  /* bad argument / shape / non-canonical input detected */
  B200SP_ERR_CUDA = -2,     /* a CUDA runtime call failed */
  B200SP_ERR_ALLOC = -3,    /* out of device or host memory (Python shim raises MemoryError) */
  B200SP_ERR_OVERFLOW = -4, /* nnz(C) or a per-row count does not fit int32 indptr */
  B200SP_ERR_STATE = -5,    /* numeric called without a matching symbolic */
  B200SP_ERR_IO = -6        /* txt-CSR file could not be read / written */
};

/* value type; mirrors CUDA_R_32F / CUDA_R_64F passed at spgemm_from_txt_alg1.cu:165
 * and cupy _dtype.to_cuda_dtype at cusparse.py:2050 */
enum b200sp_dtype { B200SP_R_32F = 0, B200SP_R_64F = 1 };

/* algorithm selector; mirrors CUSPARSE_SPGEMM_DEFAULT / ALG1 / ALG2 / ALG3 as mapped at
 * cusparse.py:2052-2057 (alg==3 -> ALG3, alg==2 -> ALG2, anything else -> DEFAULT). */
enum b200sp_alg { B200SP_ALG_DEFAULT = 0, B200SP_ALG1 = 1, B200SP_ALG2 = 2, B200SP_ALG3 = 3 };

/* ---------------------------------------------------------------- handle ---- */

int b200sp_version(void);

/* replaces cusparseCreate (spgemm_from_txt_alg1.cu:144) / device.get_cusparse_handle
 * (cusparse.py:2041).  device < 0 = current device. */
int b200sp_create(b200sp_handle_t *handle, int device);
/* replaces cusparseDestroy (spgemm_from_txt_alg1.cu:206); frees the workspace arena. */
int b200sp_destroy(b200sp_handle_t handle);
/* cudaStream_t passed as void*; NULL = legacy default stream. */
int b200sp_set_stream(b200sp_handle_t handle, void *cuda_stream);
/* Optional: route workspace allocations through the caller (the reference routes all
 * temporaries through the current CuPy allocator so its peak-VRAM sampler sees them:
 * SpGEMM_alg_comparison/profiler.py:97-98).  alloc returns a device pointer or NULL. */
int b200sp_set_allocator(b200sp_handle_t handle, void *(*alloc_fn)(size_t bytes, void *ctx),
                         void (*free_fn)(void *ptr, void *ctx), void *ctx);
/* Drop the cached workspace arena (the analogue of pool.free_all_blocks(),
 * SpGEMM_alg_comparison/profiler.py:129). */
int b200sp_release_workspace(b200sp_handle_t handle);
const char *b200sp_last_error(b200sp_handle_t handle);
const char *b200sp_status_string(int status);

/* ---------------------------------------------------------------- SpGEMM ---- */

/*
 * Symbolic phase of C = alpha * A @ B  (A: m x k, B: k x n, C: m x n).
 *
 * Replaces, in one call: cusparseSpGEMM_workEstimation x2 (spgemm_from_txt_alg1.cu:171-175,
 * cusparse.py:2061-2068), cusparseSpGEMM_getNumProducts (spgemm_from_txt_alg3.cu:190-191),
 * cusparseSpGEMM_estimateMemory x2 (spgemm_from_txt_alg3.cu:195-201, cusparse.py:2095-2105),
 * the structure half of cusparseSpGEMM_compute (alg1.cu:177-181) and cusparseSpMatGetSize
 * (alg1.cu:183).  It counts intermediate products per row, bins rows by size, computes the
 * exact nnz of every row of C and writes the final C.indptr into dCp (m+1 int32, device,
 * caller owned - allocated before the call exactly like alg1.cu:156 / cusparse.py:2039).
 *
 *   alg, chunk_fraction : mirrors cuSPARSE's ALG1 / ALG2 / ALG3 knob (chunk_fraction in (0,1],
 *       default 0.2 as in spgemm_from_txt_alg3.cu:103).  This engine never allocates an
 *       O(products) buffer: ALG1 already runs at inputs + C + 12 m bytes.  What the chunked modes
 *       can still bound is (a) here, with device-resident caller-owned C: the CTA-private spill
 *       workspace of the global-bitmap kernels (their grid shrinks by ceil(1/chunk_fraction));
 *       (b) in b200sp_spgemm_host with ALG3: C itself - see there.  Bits are identical in all modes.
 *   num_products (host, out, may be NULL): sum_i sum_{k in A_i} nnz(B_k)  (FLOPs / 2).
 *   nnzA, nnzB: stored entries of A and B (what cusparseCreateCsr is given, alg1.cu:149-154).
 *   nnzC (host, out): number of stored entries of C.  One host synchronisation happens here,
 *       as in the reference (cusparse.py:2120-2127).
 *
 * The handle keeps the plan (row bins, borrowed dAp/dAj/dBp/dBj/dCp pointers) for numeric.
 */
int b200sp_spgemm_symbolic(b200sp_handle_t handle, int64_t m, int64_t k, int64_t n, int64_t nnzA,
                           int64_t nnzB, const int32_t *dAp, const int32_t *dAj,
                           const int32_t *dBp, const int32_t *dBj, int alg, float chunk_fraction,
                           int64_t *num_products, int64_t *nnzC, int32_t *dCp);

/* OR-ed into `alg` of b200sp_spgemm_symbolic: the b200sp_spgemm_estimate that has just run on this
 * handle with exactly these arguments already did the first half (products per row, bins, canonical
 * check of B); it is not repeated.  B200SP_ERR_STATE when no such estimate is pending. */
#define B200SP_ALG_REUSE_ESTIMATE 0x100

/*
 * Single-pass SpGEMM for well-filled C, in two calls.  The reference's cuSPARSE sequence computes
 * the product into temporary buffers sized from an estimate and copies it out afterwards
 * (cusparseSpGEMM_workEstimation / _compute / _copy: spgemm_from_txt_alg1.cu:171-188,
 * cusparse.py:2061-2137); cuSPARSE's ALG1 is the variant that spends memory to save time.  This is
 * that trade here: C.indices / C.data are allocated at an UPPER BOUND of nnz(C) and the products
 * are walked once instead of twice.
 *
 * b200sp_spgemm_estimate: products per row of A @ B (replaces workEstimation + getNumProducts,
 *   spgemm_from_txt_alg3.cu:182-191) and from them
 *     *nnz_bound  = sum_i min(products_i, n) >= nnz(C)  (<= 1.6 nnz(C) for uniformly random operands),
 *     *single_pass = 1 when the single-pass kernel applies and pays: alg is not a chunked mode, an
 *       n-column accumulator of `dtype` fits shared memory, the bound fits int32, and the product
 *       count is at least a quarter of m * n (rows of C well filled; measured cross-over).
 *   dCp (m+1 int32, caller owned) is the C.indptr the following call fills.  One host sync.
 *   Then EITHER b200sp_spgemm_single_pass (only if *single_pass), OR b200sp_spgemm_symbolic with
 *   alg | B200SP_ALG_REUSE_ESTIMATE and the same arguments, followed by b200sp_spgemm_numeric.
 * b200sp_spgemm_single_pass: fills dCp, dCj, dCx (capacity >= *nnz_bound entries each) and returns
 *   the exact *nnzC; entries [nnzC, capacity) are left untouched - the caller trims (a view is
 *   enough, nothing is moved).  Rows are produced in row order by one persistent kernel: a row's
 *   offset is the total of the rows before it, obtained by decoupled look-back over per-row status
 *   words, so no separate structure pass, scan or binning runs.  Same bits as the two-pass path.
 *   One host sync (nnzC), like cusparseSpMatGetSize.
 */
int b200sp_spgemm_estimate(b200sp_handle_t handle, int dtype, int64_t m, int64_t k, int64_t n,
                           int64_t nnzA, int64_t nnzB, const int32_t *dAp, const int32_t *dAj,
                           const int32_t *dBp, const int32_t *dBj, int alg, float chunk_fraction,
                           int32_t *dCp, int64_t *num_products, int64_t *nnz_bound, int *single_pass);
int b200sp_spgemm_single_pass(b200sp_handle_t handle, int dtype, const void *alpha, const void *dAx,
                              const void *dBx, int32_t *dCj, void *dCx, int64_t capacity,
                              int64_t *nnzC);

/*
 * The single pass without the estimate's host synchronisation, for callers that can tell from the
 * sizes alone that C will be well filled (the Python operator does when nnzA * nnzB / k >= m * n / 2):
 * the caller allocates dCj / dCx at the bound that needs no estimate, capacity >= m * n entries, and
 * the products-per-row kernel, the canonical check of B, the packing of B and the single-pass kernel
 * are enqueued back to back; the kernel itself does nothing if the check of B failed.  ONE host sync
 * at the end returns *nnzC and *num_products (B200SP_ERR_INVALID for non-canonical B).  Same bits as
 * every other mode.  Requires n * sizeof(value) <= 208 KB and m * n <= 2^31 - 1.  Replaces the same
 * reference calls as b200sp_spgemm_estimate + b200sp_spgemm_single_pass.
 */
int b200sp_spgemm_single_pass_direct(b200sp_handle_t handle, int dtype, int64_t m, int64_t k, int64_t n,
                                     int64_t nnzA, int64_t nnzB, const int32_t *dAp, const int32_t *dAj,
                                     const void *dAx, const int32_t *dBp, const int32_t *dBj,
                                     const void *dBx, const void *alpha, int32_t *dCp, int32_t *dCj,
                                     void *dCx, int64_t capacity, int64_t *nnzC, int64_t *num_products);

/*
 * Numeric phase: fills dCj (nnzC int32) and dCx (nnzC values), both caller owned, allocated
 * after reading nnzC exactly like cusparse.py:2128-2131 / alg1.cu:184-186.
 * Replaces the value half of cusparseSpGEMM_compute and cusparseSpGEMM_copy (alg1.cu:187-188,
 * cusparse.py:2134-2137).  alpha is a HOST pointer to one value of `dtype`
 * (CUSPARSE_POINTER_MODE_HOST, alg1.cu:145); NULL means 1.  beta is fixed at 0 as in the
 * reference.  Columns within each row come out sorted; every column's sum is accumulated in
 * A-row storage order with a separate multiply and add (no FMA), which is bit-for-bit what
 * scipy.sparse's csr_matmat produces (SpGEMM_vs_SpMV/profiler.py:408) and is run-to-run
 * deterministic (deterministic/test_deterministic.sh).
 * May be called repeatedly with new values for the same structure.
 */
int b200sp_spgemm_numeric(b200sp_handle_t handle, int dtype, const void *alpha, const void *dAx,
                          const void *dBx, int32_t *dCj, void *dCx);

/*
 * Host-buffer one-shot: H2D of A and B, symbolic, numeric, D2H of C.  This is the whole body
 * of the reference driver between read_csr_txt_prefix and write_csr_txt_prefix
 * (spgemm_from_txt_alg1.cu:130-195) and of the profilers' timed closure
 * `to_gpu_sparse(A); to_gpu_sparse(B); cusparse.spgemm(...)` (SpGEMM_alg_comparison/profiler.py:210-213)
 * plus the .get().  The output arrays live in pinned host memory OWNED BY THE HANDLE: they stay
 * valid until the next *_host call on this handle or b200sp_destroy (do not free them).
 * hA* == hB* (aliasing) uploads the matrix once.
 * alg == B200SP_ALG3: the memory-bounded mode (spgemm_from_txt_alg3.cu:101-115,195-208).  The rows
 * of C are cut into chunks of at most chunk_fraction * nnz(C) entries (a larger row is a chunk of
 * its own); chunk c+1 is computed while chunk c is copied to the host, so the device holds two
 * chunk buffers instead of C: peak = inputs + 2 * chunk + workspace.  The cost is a second
 * structure pass per chunk and two host synchronisations per chunk.  ALG2 = ALG1 here.
 * b200sp_stats.peak_workspace_bytes after the call (on a released arena) is the device
 * high-water mark including the staged inputs and outputs; num_chunks the number of chunks.
 */
int b200sp_spgemm_host(b200sp_handle_t handle, int dtype, int64_t m, int64_t k, int64_t n,
                       const int32_t *hAp, const int32_t *hAj, const void *hAx, const int32_t *hBp,
                       const int32_t *hBj, const void *hBx, int alg, float chunk_fraction,
                       const void *alpha, int32_t **hCp, int32_t **hCj, void **hCx, int64_t *nnzC,
                       int64_t *num_products);
/* releases arrays returned by b200sp_read_csr_txt */
int b200sp_free_host(void *ptr);

/* -------------------------------------------- host-side plumbing of the row-panel driver ---- */
/* ("C panels D2H into one pinned host buffer at offsets from an exclusive
 * scan of per-panel nnzC".)  The reference has no multi-GPU path; its single-GPU drivers copy C
 * back with plain cudaMemcpy (spgemm_from_txt_alg1.cu:190-193).  These three calls let several
 * processes (one per GPU) land their C panels in ONE host buffer - e.g. a POSIX shared-memory
 * segment mapped by every rank - at DMA speed:
 *   b200sp_host_register   page-locks [ptr, ptr+bytes) for this process (cudaHostRegister,
 *                          portable), so device-to-host copies into it are asynchronous DMA;
 *   b200sp_host_unregister undoes it;
 *   b200sp_memcpy_async    cudaMemcpyAsync on a caller stream (NULL = the handle's second DMA
 *                          queue, which runs beside the kernels of the handle's compute stream).
 *                          kind: 0 = host to device, 1 = device to host. */
int b200sp_host_register(void *ptr, size_t bytes);
int b200sp_host_unregister(void *ptr);
int b200sp_memcpy_async(b200sp_handle_t handle, void *dst, const void *src, size_t bytes, int kind,
                        void *cuda_stream);

/* Host only (no device is touched): out[i] = sum over the nonzeros (i, c) of A of nnz(B_c), the
 * intermediate products of row i of A @ B, from host arrays - what the row-panel driver balances its
 * panels by before anything is uploaded (the device-side count is part of b200sp_spgemm_symbolic /
 * _estimate; cusparseSpGEMM_getNumProducts, spgemm_from_txt_alg3.cu:190-191, only gives the total).
 * One multithreaded pass (nthreads <= 0: up to 16).  B200SP_ERR_INVALID for a malformed hAp or a
 * column index outside [0, k). */
int b200sp_host_row_products(int64_t m, int64_t k, const int32_t *hAp, const int32_t *hAj,
                             const int32_t *hBp, int64_t *out, int nthreads);

/* ------------------------------------------------------------ SpMV / SpMM ---- */

/*
 * y = alpha * A @ x + beta * y, A CSR m x n.  Replaces cusparseSpMV_bufferSize + cusparseSpMV
 * (cusparse.py:1430-1435).  alpha / beta are HOST pointers to one value of `dtype` (NULL = 1 / 0).
 * When *beta == 0, y is write-only (never read), as with cuSPARSE.
 */
int b200sp_spmv(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, int64_t nnz,
                const int32_t *dAp, const int32_t *dAj, const void *dAx, const void *dx, void *dy,
                const void *alpha, const void *beta);

/*
 * C = alpha * A @ B + beta * C with dense column-major B (n x ncols, leading dimension ldb) and
 * C (m x ncols, ldc).  Replaces cusparseSpMM_bufferSize + cusparseSpMM (cusparse.py:1505-1511);
 * with ncols == 1 this is the path `A_gpu @ x_gpu` with an (n,1) vector takes in the reference
 * (cupyx/scipy/sparse/_csr.py:217-225, SpGEMM_vs_SpMV/profiler.py:490-493).
 */
int b200sp_spmm(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, int64_t nnz,
                int64_t ncols, const int32_t *dAp, const int32_t *dAj, const void *dAx,
                const void *dB, int64_t ldb, void *dC, int64_t ldc, const void *alpha,
                const void *beta);

/*
 * Allocation-free SpMV: y = alpha * A @ x + beta * y like b200sp_spmv, always with the merge-path
 * kernels, with their per-tile scratch in a fixed device array inside the library instead of the
 * handle's workspace.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot reach cudaMalloc (code under CUDA stream capture, serving
 * loops, harnesses that forbid it).  The call performs no cudaMalloc / cudaFree / allocating
 * library call; inputs and outputs are the caller's device arrays as in b200sp_spmv.
 *   limit: m + nnz <= B200SP_NOALLOC_MAX_TILES * 2048 (= 2^27) merge items; larger problems are
 *       rejected with B200SP_ERR_INVALID (message in b200sp_last_error) - use b200sp_spmv.
 *   cost: every GPU that loads the library holds the B200SP_NOALLOC_SCRATCH_BYTES array until exit.
 *   sharing: all calls on a device use the same array, which keeps the last call's contents
 *       (nothing in it needs to start at zero, every word that is read was written by the same
 *       call); calls that OVERLAP on the GPU - different streams, handles or threads - race.
 *   The first call on a device resolves the array's device address (cudaGetSymbolAddress); make
 *   one call before starting a stream capture.
 * Results are bit-identical to b200sp_spmv with the merge-path kernel (B200SP_SPMV_KERNEL=merge).
 */
#define B200SP_NOALLOC_MAX_TILES 65536
#define B200SP_NOALLOC_SCRATCH_BYTES 2359552
int b200sp_spmv_noalloc(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, int64_t nnz,
                        const int32_t *dAp, const int32_t *dAj, const void *dAx, const void *dx,
                        void *dy, const void *alpha, const void *beta);

/* Host-buffer one-shot SpMV (H2D of A and x, kernel, D2H of y): the profilers' timed closure
 * `to_gpu_sparse(A); cp.asarray(C); A_gpu @ C_gpu` (SpGEMM_vs_SpMV/profiler.py:490-493). */
int b200sp_spmv_host(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, int64_t nnz,
                     const int32_t *hAp, const int32_t *hAj, const void *hAx, const void *hx,
                     void *hy, const void *alpha, const void *beta);

/* -------------------------------------------------- canonical-format helpers ---- */

/* 1 in *result when every row of the CSR has strictly increasing column indices (sorted, no
 * duplicates) and all indices are in [0, ncols).  Replaces the canonical-format elementwise
 * kernel + .all() (cupyx/scipy/sparse/_compressed.py:177-192) behind has_canonical_format. */
int b200sp_csr_is_canonical(b200sp_handle_t handle, int64_t m, int64_t ncols, const int32_t *dAp,
                            const int32_t *dAj, int *result);

/* In-row column sort of a CSR (values follow their columns).  Replaces cusparseXcsrsort +
 * gather used by csr_matrix.sort_indices() (cupyx/cusparse.py:832-866). */
int b200sp_csr_sort_indices(b200sp_handle_t handle, int dtype, int64_t m, int64_t nnz,
                            const int32_t *dAp, int32_t *dAj, void *dAx);

/* ------------------------------------- format conversion and canonicalisation ---- */
/*   All arrays are device pointers; int32 indices, base 0.  Scratch
 * (~24.3 bytes per entry) comes from the handle's workspace.  Every call synchronises once. */

/* COO -> canonical CSR: entries sorted by (row, column), duplicates summed.  Duplicates are
 * added sequentially in INPUT order with separately rounded adds (deterministic; the reference
 * uses floating-point atomics).  dCp has m+1 entries; dCj/dCx need room for nnz entries, of
 * which *nnz_out (host) are written.  Replaces coo_matrix.sum_duplicates + tocsr
 * (cupyx/scipy/sparse/_coo.py:356-459, 517-540; cupyx/cusparse.py coosort :908-955,
 * coo2csr :957-970).  B200SP_ERR_INVALID for a coordinate outside [0,m) x [0,n). */
int b200sp_coo_to_csr(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, int64_t nnz,
                      const int32_t *dRow, const int32_t *dCol, const void *dVal, int32_t *dCp,
                      int32_t *dCj, void *dCx, int64_t *nnz_out);

/* CSR with unsorted rows and/or duplicate entries -> canonical CSR (out of place), same summation
 * rule.  Replaces csr_matrix.sum_duplicates (cupyx/scipy/sparse/_compressed.py:971-991). */
int b200sp_csr_sum_duplicates(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, int64_t nnz,
                              const int32_t *dAp, const int32_t *dAj, const void *dAx,
                              int32_t *dCp, int32_t *dCj, void *dCx, int64_t *nnz_out);

/* CSC (n+1 column pointers, row indices) of an m x n matrix -> CSR with sorted rows; entries are
 * only re-ordered (nnz is preserved, duplicates stay adjacent in input order).  Replaces
 * csc_matrix.tocsr (cupyx/scipy/sparse/_csc.py:304-326; cupyx/cusparse.py csc2csr :1092-1114). */
int b200sp_csc_to_csr(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, int64_t nnz,
                      const int32_t *dColPtr, const int32_t *dRowIdx, const void *dVal,
                      int32_t *dCp, int32_t *dCj, void *dCx);

/* CSR of an m x n matrix -> CSC (equivalently: the CSR of the transpose).  Replaces
 * csr_matrix.tocsc / .T (cupyx/scipy/sparse/_csr.py:455-477; cupyx/cusparse.py csr2csc :1014-1036). */
int b200sp_csr_to_csc(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, int64_t nnz,
                      const int32_t *dAp, const int32_t *dAj, const void *dAx, int32_t *dColPtr,
                      int32_t *dRowIdx, void *dVal);

/* ------------------------------------------------ dense image and comparison ---- */
/*  */

#define B200SP_ORDER_ROW 0 /* C order: out[i * n + j] */
#define B200SP_ORDER_COL 1 /* F order: out[i + j * m] */

/* CSR -> dense m x n array (zero-filled first; duplicates accumulate in storage order).
 * Replaces csr_matrix.toarray (cupyx/scipy/sparse/_csr.py:383-420; cusparse.csr2dense :768-798). */
int b200sp_csr_to_dense(b200sp_handle_t handle, int dtype, int64_t m, int64_t n, const int32_t *dAp,
                        const int32_t *dAj, const void *dAx, void *dOut, int order);

/* max |A - B| over all positions of two canonical CSR matrices with m rows (an entry stored in
 * only one of them is compared with 0); NaN if any compared value is NaN.  The fused form of
 * `abs(C.toarray() - D.toarray()).max()` (numerical_error/error.py:25-29) without the two
 * m x n dense arrays.  *result is a host double. */
int b200sp_csr_max_abs_diff(b200sp_handle_t handle, int dtype, int64_t m, const int32_t *dAp,
                            const int32_t *dAj, const void *dAx, const int32_t *dBp,
                            const int32_t *dBj, const void *dBx, double *result);

/* ----------------------------------------------------------- statistics ---- */

#define B200SP_MAX_BINS 16

typedef struct b200sp_stats {
  int64_t num_products;       /* of the last symbolic */
  int64_t nnz_c;              /* of the last symbolic */
  int64_t workspace_bytes;    /* arena currently held by the handle */
  int64_t peak_workspace_bytes; /* high-water mark of the arena since create / reset */
  int64_t max_row_products;   /* largest per-row intermediate-product count */
  int32_t num_chunks;         /* row chunks used by the last symbolic (1 unless ALG2/ALG3) */
  int32_t num_bins;
  int64_t sym_bin_rows[B200SP_MAX_BINS]; /* rows per symbolic bin (by product upper bound) */
  int64_t num_bin_rows[B200SP_MAX_BINS]; /* rows per numeric bin (by exact nnz of the C row) */
  int64_t kernel_launches;    /* kernels launched by this handle since create / reset */
  float ms_symbolic;          /* device time of the last symbolic / numeric (CUDA events) */
  float ms_numeric;
} b200sp_stats;

int b200sp_get_stats(b200sp_handle_t handle, b200sp_stats *stats);
int b200sp_reset_stats(b200sp_handle_t handle);

/* ------------------------------------------------------------- txt-CSR I/O ---- */

/*
 * The reference's on-disk format: <prefix>_indptr.txt, <prefix>_indices.txt, <prefix>_data.txt,
 * one number per line (np.savetxt "%d" / "%.9g", cupy_cusparse/gen_and_save_alg1_txt.py:12-14).
 * Reader replaces read_csr_txt_prefix (spgemm_from_txt_alg1.cu:19-34,52-70): integers parsed as
 * long long then narrowed to int32, values parsed as double then narrowed to float.  rows =
 * len(indptr)-1, nnz = len(indices); the column count is not stored in the format.
 * Arrays are malloc'ed by the library; release with b200sp_free_host.
 */
int b200sp_read_csr_txt(const char *prefix, int64_t *rows, int64_t *nnz, int32_t **hp, int32_t **hj,
                        float **hx, char *errbuf, size_t errbuf_len);
/* Writer replaces write_csr_txt_prefix (spgemm_from_txt_alg1.cu:35-46,72-76): values printed
 * with 9 significant digits in the default (%g-like) float format, which round-trips fp32. */
int b200sp_write_csr_txt(const char *prefix, int64_t rows, int64_t nnz, const int32_t *hp,
                         const int32_t *hj, const float *hx, char *errbuf, size_t errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* B200SP_H_ */
