// gather_probe.cu -- how fast can an SM do random 4-byte gathers from an L2-resident vector?
// (sets the ceiling of CSR SpMV with uniformly random columns: one gather per nonzero).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

template <int U>
__global__ void k_gather(const int *__restrict__ idx, const float *__restrict__ x, float *__restrict__ out, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride * U) {
    int j[U];
#pragma unroll
    for (int u = 0; u < U; ++u) j[u] = (i + u * stride < n) ? __ldcs(idx + i + u * stride) : 0;
#pragma unroll
    for (int u = 0; u < U; ++u) acc += __ldg(x + j[u]);
  }
  out[blockIdx.x * (long long)blockDim.x + threadIdx.x] = acc;
}

template <int U>
__global__ void k_gather_mul(const int *__restrict__ idx, const float *__restrict__ val, const float *__restrict__ x,
                             float *__restrict__ out, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride * U) {
    int j[U];
    float v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const bool ok = i + u * stride < n;
      j[u] = ok ? __ldcs(idx + i + u * stride) : 0;
      v[u] = ok ? __ldcs(val + i + u * stride) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) acc += v[u] * __ldg(x + j[u]);
  }
  out[blockIdx.x * (long long)blockDim.x + threadIdx.x] = acc;
}

int main() {
  const long long n = 11000000;
  for (int logx = 16; logx <= 24; logx += 2) {
    const int nx = 1 << logx;
    std::vector<int> h(n);
    unsigned long long s = 88172645463325252ULL;
    for (long long i = 0; i < n; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int)(s % nx); }
    int *idx; float *x, *out, *val;
    cudaMalloc(&val, n * 4); cudaMemset(val, 0, n * 4);
    cudaMalloc(&idx, n * 4); cudaMalloc(&x, (size_t)nx * 4); cudaMalloc(&out, 148 * 16 * 256 * 4);
    cudaMemcpy(idx, h.data(), n * 4, cudaMemcpyHostToDevice);
    cudaMemset(x, 0, (size_t)nx * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int cta = 4; cta <= 8; cta += 4) {
      float best = 1e9f;
      for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        k_gather<8><<<148 * cta, 256>>>(idx, x, out, n);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
      }
      float best2 = 1e9f;
      static char *fl = nullptr;
      if (!fl) cudaMalloc(&fl, 256 << 20);
      for (int rep = 0; rep < 6; ++rep) {
        cudaMemsetAsync(fl, rep, 256 << 20);
        cudaEventRecord(e0);
        k_gather_mul<8><<<148 * cta, 256>>>(idx, val, x, out, n);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best2) best2 = ms;
      }
      printf("   with a streamed value array, L2 flushed before each run: %7.1f us\n", best2 * 1e3);
      printf("x=%8d floats (%6.1f MB) ctas/SM=%d : %7.1f us  -> %.2f SM-cycles/gather @1.965GHz, %.0f Ggather/s\n", nx, nx * 4 / 1e6, cta,
             best * 1e3, best * 1e-3 * 1.965e9 * 148 / n, n / best / 1e6);
    }
    cudaFree(val); cudaFree(idx); cudaFree(x); cudaFree(out);
  }
  return 0;
}
