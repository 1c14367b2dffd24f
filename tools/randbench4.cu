// randbench4.cu -- is the 36.8 G/s random-read ceiling DRAM or address translation?  Footprint sweep (64 MiB .. 64 GiB)
// of independent random 8-byte reads (ld.global.L2::64B) and writes, 65,536 threads x K = 16 in flight.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o randbench4.bin randbench4.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

template <int WRITE, int K>
__global__ void k_acc(double *tbl, uint64_t mask, int iters, double *sink)
{
  uint64_t s = (uint64_t)(blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345;
  double acc = 0;
  for (int it = 0; it < iters; ++it)
  {
    uint64_t ix[K];
#pragma unroll
    for (int k = 0; k < K; ++k)
    {
      s = s * 6364136223846793005ull + 1442695040888963407ull;
      ix[k] = (s >> 17) & mask;
    }
    if (!WRITE)
    {
      double v[K];
#pragma unroll
      for (int k = 0; k < K; ++k) asm volatile("ld.global.L2::64B.f64 %0, [%1];" : "=d"(v[k]) : "l"(tbl + ix[k]));
#pragma unroll
      for (int k = 0; k < K; ++k) acc += v[k];
    }
    else
    {
#pragma unroll
      for (int k = 0; k < K; ++k) tbl[ix[k]] = (double)it;
    }
  }
  if (acc == 123.456) *sink = acc;
}

template <int WRITE>
static void run(double *tbl, uint64_t elems, int threads, double *sink)
{
  const int iters = 64, K = 16;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_acc<WRITE, K><<<threads / 64, 64>>>(tbl, elems - 1, 2, sink);
  cudaEventRecord(e0);
  k_acc<WRITE, K><<<threads / 64, 64>>>(tbl, elems - 1, iters, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  double acc = (double)threads * iters * K;
  printf("footprint %7llu MiB  threads %6d  %s  %7.2f ms  %7.2f G accesses/s   (%s)\n", (unsigned long long)(elems * 8 >> 20), threads,
         WRITE ? "write" : "read ", ms, acc / ms * 1e-6, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
  double *sink, *tbl;
  cudaMalloc(&sink, 8);
  const uint64_t max_elems = 1ull << 33;       // 64 GiB
  if (cudaMalloc(&tbl, max_elems * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(tbl, 0, max_elems * 8);
  for (uint64_t elems = 1ull << 23; elems <= max_elems; elems <<= 1)
  {
    run<0>(tbl, elems, 65536, sink);
    run<1>(tbl, elems, 65536, sink);
  }
  run<0>(tbl, 1ull << 27, 16384, sink);        // 1 GiB, fewer threads
  run<0>(tbl, 1ull << 27, 262144, sink);
  cudaDeviceSynchronize();
  return 0;
}
