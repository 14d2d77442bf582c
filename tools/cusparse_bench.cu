// cusparse_bench.cu -- COMPARATOR, not product code: times cuSPARSE SpGEMM ALG1/ALG2/ALG3 with
// device-resident inputs, issuing exactly the call sequence of the reference drivers
// (cupy_cusparse/spgemm_from_txt_alg{1,3}.cu:171-218), and records the peak of live device
// allocations (inputs excluded and included), the figure the reference samples with
// cudaMemGetInfo (SpGEMM_alg_comparison/profiler.py:82-143).
//
//   cusparse_bench <A.bin> <alg 1|2|3> [chunk_fraction] [reps]
//   cusparse_bench <A.bin> 0 [unused] [reps]      cusparseSpMV y = A x (CUSPARSE_SPMV_ALG_DEFAULT and CSR_ALG2),
//                                                 L2 flushed before every timed run
// A.bin: int64 m, int64 n, int64 nnz, int32 indptr[m+1], int32 indices[nnz], float data[nnz]
// Computes A @ A.  Prints one JSON line.
#include <cuda_runtime_api.h>
#include <cusparse.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t s = (x); if (s != cudaSuccess) { printf("{\"error\": \"cuda %s line %d\"}\n", cudaGetErrorString(s), __LINE__); return 1; } } while (0)
#define CS(x) do { cusparseStatus_t s = (x); if (s != CUSPARSE_STATUS_SUCCESS) { printf("{\"error\": \"cusparse status %d line %d\"}\n", (int)s, __LINE__); return 1; } } while (0)

static size_t g_live = 0, g_peak = 0;
static cudaError_t tmalloc(void **p, size_t b) {
  cudaError_t e = cudaMalloc(p, b);
  if (e == cudaSuccess) { g_live += b; g_peak = std::max(g_peak, g_live); }
  return e;
}
static void tfree(void *p, size_t b) { if (p) { cudaFree(p); g_live -= b; } }

int main(int argc, char **argv) {
  if (argc < 3) { fprintf(stderr, "usage: %s A.bin alg [chunk_fraction] [reps]\n", argv[0]); return 2; }
  const int algn = atoi(argv[2]);
  const float cf = argc > 3 ? (float)atof(argv[3]) : 0.2f;
  const int reps = argc > 4 ? atoi(argv[4]) : 5;
  FILE *f = fopen(argv[1], "rb");
  if (!f) { perror("open"); return 1; }
  int64_t hdr[3];
  if (fread(hdr, 8, 3, f) != 3) return 1;
  const int64_t m = hdr[0], n = hdr[1], nnz = hdr[2];
  std::vector<int32_t> hp(m + 1), hj(nnz);
  std::vector<float> hx(nnz);
  if (fread(hp.data(), 4, m + 1, f) != (size_t)(m + 1) || fread(hj.data(), 4, nnz, f) != (size_t)nnz ||
      fread(hx.data(), 4, nnz, f) != (size_t)nnz) return 1;
  fclose(f);
  int32_t *dp, *dj, *dCp;
  float *dx;
  CK(tmalloc((void **)&dp, (m + 1) * 4)); CK(tmalloc((void **)&dj, nnz * 4)); CK(tmalloc((void **)&dx, nnz * 4));
  CK(cudaMemcpy(dp, hp.data(), (m + 1) * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dj, hj.data(), nnz * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dx, hx.data(), nnz * 4, cudaMemcpyHostToDevice));
  const size_t input_bytes = g_live;
  cusparseHandle_t h; CS(cusparseCreate(&h));
  CS(cusparseSetPointerMode(h, CUSPARSE_POINTER_MODE_HOST));
  if (algn == 0) {  // ---- SpMV comparator (the call cupyx.cusparse.spmv makes, cusparse.py:1430-1435)
    float *dxv, *dy; char *flush;
    CK(cudaMalloc((void **)&dxv, n * 4)); CK(cudaMalloc((void **)&dy, m * 4)); CK(cudaMalloc((void **)&flush, 512 << 20));
    std::vector<float> hxv(n);
    for (int64_t i = 0; i < n; ++i) hxv[i] = (float)((i * 2654435761u) % 1000) / 1000.f - 0.5f;
    CK(cudaMemcpy(dxv, hxv.data(), n * 4, cudaMemcpyHostToDevice));
    cusparseSpMatDescr_t A; cusparseDnVecDescr_t X, Y;
    CS(cusparseCreateCsr(&A, m, n, nnz, dp, dj, dx, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F));
    CS(cusparseCreateDnVec(&X, n, dxv, CUDA_R_32F)); CS(cusparseCreateDnVec(&Y, m, dy, CUDA_R_32F));
    const float al = 1.f, be = 0.f;
    cudaEvent_t s0, s1; cudaEventCreate(&s0); cudaEventCreate(&s1);
    const cusparseSpMVAlg_t algs[2] = {CUSPARSE_SPMV_ALG_DEFAULT, CUSPARSE_SPMV_CSR_ALG2};
    const char *names[2] = {"default", "csr_alg2"};
    printf("{\"impl\": \"cusparse\", \"op\": \"spmv\", \"m\": %lld, \"nnz\": %lld", (long long)m, (long long)nnz);
    for (int a = 0; a < 2; ++a) {
      size_t bs = 0; void *buf = nullptr;
      CS(cusparseSpMV_bufferSize(h, CUSPARSE_OPERATION_NON_TRANSPOSE, &al, A, X, &be, Y, CUDA_R_32F, algs[a], &bs));
      if (bs) CK(cudaMalloc(&buf, bs));
      std::vector<float> ts;
      for (int it = 0; it < reps + 2; ++it) {
        CK(cudaMemsetAsync(flush, it, 512 << 20));
        cudaEventRecord(s0);
        CS(cusparseSpMV(h, CUSPARSE_OPERATION_NON_TRANSPOSE, &al, A, X, &be, Y, CUDA_R_32F, algs[a], buf));
        cudaEventRecord(s1);
        CK(cudaEventSynchronize(s1));
        float ms; cudaEventElapsedTime(&ms, s0, s1);
        if (it >= 2) ts.push_back(ms);
      }
      std::sort(ts.begin(), ts.end());
      printf(", \"ms_%s\": %.4f, \"buffer_%s\": %zu", names[a], ts[ts.size() / 2], names[a], bs);
      if (buf) cudaFree(buf);
    }
    printf("}\n");
    return 0;
  }
  const cusparseSpGEMMAlg_t alg = algn == 3 ? CUSPARSE_SPGEMM_ALG3 : algn == 2 ? CUSPARSE_SPGEMM_ALG2 : CUSPARSE_SPGEMM_ALG1;
  const float alpha = 1.f, beta = 0.f;
  const cusparseOperation_t op = CUSPARSE_OPERATION_NON_TRANSPOSE;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  std::vector<float> times;
  int64_t Cnnz = 0, prods = 0;
  size_t peak_excl = 0;
  for (int it = 0; it < reps + 1; ++it) {  // first iteration is the warm-up
    g_peak = g_live;
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    cusparseSpMatDescr_t A, B, C;
    CS(cusparseCreateCsr(&A, m, n, nnz, dp, dj, dx, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F));
    CS(cusparseCreateCsr(&B, m, n, nnz, dp, dj, dx, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F));
    CK(tmalloc((void **)&dCp, (m + 1) * 4));
    CS(cusparseCreateCsr(&C, m, n, 0, dCp, nullptr, nullptr, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_32I, CUSPARSE_INDEX_BASE_ZERO, CUDA_R_32F));
    cusparseSpGEMMDescr_t d; CS(cusparseSpGEMM_createDescr(&d));
    size_t b1 = 0, b2 = 0, b3 = 0; void *B1 = nullptr, *B2 = nullptr, *B3 = nullptr;
    CS(cusparseSpGEMM_workEstimation(h, op, op, &alpha, A, B, &beta, C, CUDA_R_32F, alg, d, &b1, nullptr));
    if (b1) CK(tmalloc(&B1, b1));
    CS(cusparseSpGEMM_workEstimation(h, op, op, &alpha, A, B, &beta, C, CUDA_R_32F, alg, d, &b1, B1));
    if (algn == 1) {
      CS(cusparseSpGEMM_compute(h, op, op, &alpha, A, B, &beta, C, CUDA_R_32F, alg, d, &b2, nullptr));
      if (b2) CK(tmalloc(&B2, b2));
      CS(cusparseSpGEMM_compute(h, op, op, &alpha, A, B, &beta, C, CUDA_R_32F, alg, d, &b2, B2));
    } else {
      CS(cusparseSpGEMM_getNumProducts(d, &prods));
      CS(cusparseSpGEMM_estimateMemory(h, op, op, &alpha, A, B, &beta, C, CUDA_R_32F, alg, d, cf, &b3, nullptr, nullptr));
      if (b3) CK(tmalloc(&B3, b3));
      CS(cusparseSpGEMM_estimateMemory(h, op, op, &alpha, A, B, &beta, C, CUDA_R_32F, alg, d, cf, &b3, B3, &b2));
      tfree(B3, b3); B3 = nullptr;
      if (b2) CK(tmalloc(&B2, b2));
      CS(cusparseSpGEMM_compute(h, op, op, &alpha, A, B, &beta, C, CUDA_R_32F, alg, d, &b2, B2));
    }
    int64_t Cr, Cc;
    CS(cusparseSpMatGetSize(C, &Cr, &Cc, &Cnnz));
    int32_t *dCj; float *dCx;
    CK(tmalloc((void **)&dCj, std::max<int64_t>(Cnnz, 1) * 4)); CK(tmalloc((void **)&dCx, std::max<int64_t>(Cnnz, 1) * 4));
    CS(cusparseCsrSetPointers(C, dCp, dCj, dCx));
    CS(cusparseSpGEMM_copy(h, op, op, &alpha, A, B, &beta, C, CUDA_R_32F, alg, d));
    cudaEventRecord(e1);
    CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (it > 0) times.push_back(ms);
    peak_excl = g_peak - input_bytes;
    tfree(B1, b1); tfree(B2, b2);
    tfree(dCj, std::max<int64_t>(Cnnz, 1) * 4); tfree(dCx, std::max<int64_t>(Cnnz, 1) * 4); tfree(dCp, (m + 1) * 4);
    cusparseSpGEMM_destroyDescr(d);
    cusparseDestroySpMat(A); cusparseDestroySpMat(B); cusparseDestroySpMat(C);
  }
  std::sort(times.begin(), times.end());
  printf("{\"impl\": \"cusparse\", \"alg\": %d, \"chunk_fraction\": %g, \"m\": %lld, \"nnzA\": %lld, \"nnzC\": %lld, "
         "\"ms_median\": %.4f, \"ms_min\": %.4f, \"peak_bytes_excl_inputs\": %zu, \"input_bytes\": %zu, \"reps\": %d}\n",
         algn, cf, (long long)m, (long long)nnz, (long long)Cnnz, times[times.size() / 2], times[0], peak_excl,
         input_bytes, reps);
  return 0;
}
