// randbench2.cu -- what does ONE random table access cost in DRAM traffic on B200?
// Modes (16 GiB footprint, far larger than L2), each thread issues K independent accesses per iteration:
//   0 read 8 B at a random address            1 read 8 B in 2 sectors of one random 128 B line
//   2 read 8 B in all 4 sectors of one line   3 write 8 B at a random address (partial sector)
//   4 write one full 32 B sector              5 read-modify-write 8 B
//   6 write one full 128 B line (4 x 32 B)
// Run plain for accesses/s, and under `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,...`
// for bytes per access.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o randbench2 randbench2.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

template <int MODE, int K>
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
    if (MODE == 0 || MODE == 5)
    {
      double v[K];
#pragma unroll
      for (int k = 0; k < K; ++k) v[k] = tbl[ix[k]];
#pragma unroll
      for (int k = 0; k < K; ++k) { if (MODE == 5) tbl[ix[k]] = v[k] + 1.0; acc += v[k]; }
    }
    else if (MODE == 1 || MODE == 2)
    {
      constexpr int NS = MODE == 1 ? 2 : 4;
      double v[K][NS];
#pragma unroll
      for (int k = 0; k < K; ++k)
#pragma unroll
        for (int q = 0; q < NS; ++q) v[k][q] = tbl[(ix[k] & ~15ull) + 4 * q];
#pragma unroll
      for (int k = 0; k < K; ++k)
#pragma unroll
        for (int q = 0; q < NS; ++q) acc += v[k][q];
    }
    else if (MODE == 3)
    {
#pragma unroll
      for (int k = 0; k < K; ++k) tbl[ix[k]] = (double)it;
    }
    else if (MODE == 4 || MODE == 6)
    {
      constexpr int NS = MODE == 4 ? 1 : 4;
#pragma unroll
      for (int k = 0; k < K; ++k)
      {
        double2 *p = reinterpret_cast<double2 *>(tbl + (MODE == 4 ? (ix[k] & ~3ull) : (ix[k] & ~15ull)));
#pragma unroll
        for (int q = 0; q < 2 * NS; ++q) p[q] = make_double2((double)it, 1.0);
      }
    }
  }
  if (acc == 123.456) *sink = acc;
}

template <int MODE, int K>
static void run(double *tbl, uint64_t elems, int threads, int iters, double *sink)
{
  static const char *names[] = {"read8", "read8x2-in-line", "read8x4-in-line", "write8", "write32", "rmw8", "write128"};
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int block = 64, grid = threads / block;
  cudaEventRecord(e0);
  k_acc<MODE, K><<<grid, block>>>(tbl, elems - 1, iters, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  double acc = (double)threads * iters * K;
  printf("mode=%d %-16s threads=%7d K=%2d accesses=%.0f : %7.2f ms  %8.2f G accesses/s\n", MODE, names[MODE], threads, K,
         acc, ms, acc / ms * 1e-6);
}

int main(int argc, char **argv)
{
  int fetch = argc > 1 ? atoi(argv[1]) : 32;
  cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, fetch);
  size_t got = 0;
  cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
  printf("L2 fetch granularity requested %d, reported %zu\n", fetch, got);
  double *sink;
  cudaMalloc(&sink, 8);
  const uint64_t elems = 1ull << 31;   // 16 GiB
  double *tbl;
  if (cudaMalloc(&tbl, elems * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(tbl, 0, elems * 8);
  const int threads = 131072, iters = 64;      // 131072*64*16 = 134 M accesses per launch
  run<0, 16>(tbl, elems, threads, iters, sink);
  run<1, 16>(tbl, elems, threads, iters, sink);
  run<2, 8>(tbl, elems, threads, iters * 2, sink);
  run<3, 16>(tbl, elems, threads, iters, sink);
  run<4, 16>(tbl, elems, threads, iters, sink);
  run<5, 16>(tbl, elems, threads, iters, sink);
  run<6, 8>(tbl, elems, threads, iters * 2, sink);
  cudaDeviceSynchronize();
  return 0;
}
