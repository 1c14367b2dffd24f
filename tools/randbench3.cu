// randbench3.cu -- does any load/store flavour make a random 8-byte table access cheaper on B200?
// 16 GiB footprint, 131072 threads x K=16 independent accesses per iteration.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o randbench3.bin randbench3.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

enum { LD_PLAIN, LD_NC, LD_CG, LD_CS, LD_LU, LD_CV, LD_L2_64, LD_NOALLOC, LD_L2_64_NOALLOC, ST_PLAIN, ST_CG, ST_CS, ST_WT, RED_ADD, RMW_PLAIN, NMODES };
static const char *names[] = {"ld", "ld.nc", "ld.cg", "ld.cs", "ld.lu", "ld.cv", "ld.L2::64B", "ld.L1::no_allocate", "ld.L1::no_allocate.L2::64B",
                              "st", "st.cg", "st.cs", "st.wt", "red.add.f64", "ld+add+st"};

template <int MODE> __device__ __forceinline__ double ld(const double *p)
{
  double v;
  if (MODE == LD_PLAIN) v = *p;
  else if (MODE == LD_NC) v = __ldg(p);
  else if (MODE == LD_CG) v = __ldcg(p);
  else if (MODE == LD_CS) v = __ldcs(p);
  else if (MODE == LD_LU) v = __ldlu(p);
  else if (MODE == LD_CV) v = __ldcv(p);
  else if (MODE == LD_L2_64) asm volatile("ld.global.L2::64B.f64 %0, [%1];" : "=d"(v) : "l"(p));
  else if (MODE == LD_NOALLOC) asm volatile("ld.global.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  else asm volatile("ld.global.L1::no_allocate.L2::64B.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}

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
    if (MODE < ST_PLAIN)
    {
      double v[K];
#pragma unroll
      for (int k = 0; k < K; ++k) v[k] = ld<MODE>(tbl + ix[k]);
#pragma unroll
      for (int k = 0; k < K; ++k) acc += v[k];
    }
    else if (MODE == RMW_PLAIN)
    {
      double v[K];
#pragma unroll
      for (int k = 0; k < K; ++k) v[k] = tbl[ix[k]];
#pragma unroll
      for (int k = 0; k < K; ++k) tbl[ix[k]] = v[k] + 1.0;
    }
    else
    {
#pragma unroll
      for (int k = 0; k < K; ++k)
      {
        double *p = tbl + ix[k];
        const double v = (double)it;
        if (MODE == ST_PLAIN) *p = v;
        else if (MODE == ST_CG) __stcg(p, v);
        else if (MODE == ST_CS) __stcs(p, v);
        else if (MODE == ST_WT) __stwt(p, v);
        else asm volatile("red.global.add.f64 [%0], %1;" :: "l"(p), "d"(1.0) : "memory");
      }
    }
  }
  if (acc == 123.456) *sink = acc;
}

template <int MODE>
static void run(double *tbl, uint64_t elems, double *sink)
{
  const int threads = 131072, iters = 64, K = 16;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_acc<MODE, K><<<threads / 64, 64>>>(tbl, elems - 1, 2, sink);
  cudaEventRecord(e0);
  k_acc<MODE, K><<<threads / 64, 64>>>(tbl, elems - 1, iters, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  double acc = (double)threads * iters * K;
  printf("%-28s %7.2f ms  %7.2f G accesses/s   (%s)\n", names[MODE], ms, acc / ms * 1e-6, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
  double *sink, *tbl;
  cudaMalloc(&sink, 8);
  const uint64_t elems = 1ull << 31;
  if (cudaMalloc(&tbl, elems * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(tbl, 0, elems * 8);
  run<LD_PLAIN>(tbl, elems, sink); run<LD_NC>(tbl, elems, sink); run<LD_CG>(tbl, elems, sink); run<LD_CS>(tbl, elems, sink);
  run<LD_LU>(tbl, elems, sink); run<LD_CV>(tbl, elems, sink); run<LD_L2_64>(tbl, elems, sink); run<LD_NOALLOC>(tbl, elems, sink);
  run<LD_L2_64_NOALLOC>(tbl, elems, sink);
  run<ST_PLAIN>(tbl, elems, sink); run<ST_CG>(tbl, elems, sink); run<ST_CS>(tbl, elems, sink); run<ST_WT>(tbl, elems, sink);
  run<RED_ADD>(tbl, elems, sink); run<RMW_PLAIN>(tbl, elems, sink);
  cudaDeviceSynchronize();
  return 0;
}
