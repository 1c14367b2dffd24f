// randbench.cu -- microbenchmark of the hardware ceiling for the table access pattern of the GRL
// path: independent random 8-byte gathers (and read-modify-write scatters) over a large FP64 array.
// Each access touches its own 32-byte sector; this measures sectors/s as a function of footprint,
// threads and loads in flight per thread.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

template <int K, bool RMW>
__global__ void k_rand(double *tbl, uint64_t mask, int iters, double *sink)
{
  uint64_t s = (uint64_t)(blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345;
  double acc = 0;
  for (int it = 0; it < iters; ++it)
  {
    uint64_t ix[K];
    double v[K];
#pragma unroll
    for (int k = 0; k < K; ++k)
    {
      s = s * 6364136223846793005ull + 1442695040888963407ull;
      ix[k] = (s >> 17) & mask;
    }
#pragma unroll
    for (int k = 0; k < K; ++k) v[k] = tbl[ix[k]];
#pragma unroll
    for (int k = 0; k < K; ++k)
    {
      if (RMW) tbl[ix[k]] = v[k] + 1.0;
      acc += v[k];
    }
  }
  if (acc == 123.456) *sink = acc;
}

template <int K, bool RMW>
static void run(double *tbl, uint64_t elems, int threads, int iters, double *sink, int fetch)
{
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int block = 64, grid = threads / block;
  k_rand<K, RMW><<<grid, block>>>(tbl, elems - 1, 2, sink);
  cudaEventRecord(e0);
  k_rand<K, RMW><<<grid, block>>>(tbl, elems - 1, iters, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  double acc = (double)threads * iters * K;
  printf("fetch=%d footprint=%8.0f MiB threads=%7d K=%2d %s: %7.2f ms  %8.2f G accesses/s  (%7.1f GB/s of 32B sectors%s)\n",
         fetch, elems * 8.0 / 1048576, threads, K, RMW ? "rmw " : "read", ms, acc / ms * 1e-6,
         acc / ms * 1e-6 * 32 * (RMW ? 2 : 1), RMW ? ", rd+wr" : "");
}

int main(int argc, char **argv)
{
  int fetch = argc > 1 ? atoi(argv[1]) : 32;
  cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, fetch);
  double *sink;
  cudaMalloc(&sink, 8);
  const uint64_t max_elems = 1ull << 32;   // 32 GiB
  double *tbl;
  if (cudaMalloc(&tbl, max_elems * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(tbl, 0, max_elems * 8);
  for (uint64_t elems : {1ull << 22, 1ull << 26, 1ull << 30, 1ull << 32})
    for (int threads : {16384, 65536, 262144})
    {
      int iters = (int)(200000000ull / ((uint64_t)threads * 16));
      run<16, false>(tbl, elems, threads, iters, sink, fetch);
      run<16, true>(tbl, elems, threads, iters, sink, fetch);
      if (threads == 65536)
      {
        run<4, false>(tbl, elems, threads, iters * 4, sink, fetch);
        run<48, false>(tbl, elems, threads, iters / 3, sink, fetch);
      }
    }
  return 0;
}
