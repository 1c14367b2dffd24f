// randbench6.cu -- what would the next cached-update design buy?  A model of one agent-step's memory operations:
// per thread and iteration
//   R  random 8-byte reads (ld.global.L2::64B) over a 16 GiB table           (the Q gathers)
//   H  8-byte stores to H of the words just read                               (directory sentinels: L2 hits, write-back only)
//   M  8-byte stores to random words that were NOT read in this iteration      (write-backs of popped entries and, in the
//                                                                               "pending" design without stable slots, the
//                                                                               sentinel rewrites: fill + write-back)
//   KS coalesced 8-byte read+write stream elements over a 256 MiB buffer        (the dense trace value cache)
// Mixes: today's kernel (R 64, H 16, M 14), "p registered at act time" with and without stable slots, and the
// compulsory-traffic floor.  See DESIGN.md section 5 ("Next kernel step").
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o randbench6.bin randbench6.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

template <int R, int H, int M, int KS>
__global__ void k_mix(double *tbl, uint64_t tmask, double *stream, uint64_t smask, int iters, double *sink)
{
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
  uint64_t s = (uint64_t)tid * 0x9E3779B97F4A7C15ull + 12345;
  double acc = 0;
  for (int it = 0; it < iters; ++it)
  {
    uint64_t ix[R], mx[M > 0 ? M : 1];
#pragma unroll
    for (int k = 0; k < R; ++k) { s = s * 6364136223846793005ull + 1442695040888963407ull; ix[k] = (s >> 17) & tmask; }
#pragma unroll
    for (int k = 0; k < M; ++k) { s = s * 6364136223846793005ull + 1442695040888963407ull; mx[k] = (s >> 17) & tmask; }
    double v[R];
#pragma unroll
    for (int k = 0; k < R; ++k) asm volatile("ld.global.L2::64B.f64 %0, [%1];" : "=d"(v[k]) : "l"(tbl + ix[k]));
#pragma unroll
    for (int k = 0; k < R; ++k) acc += v[k];
    // stores to words fetched a moment ago (the value depends on the loads, as a sentinel store follows the gather)
#pragma unroll
    for (int k = 0; k < H; ++k) tbl[ix[k * (R / (H > 0 ? H : 1))]] = acc + k;
    // the dense stream
#pragma unroll 8
    for (int k = 0; k < KS; ++k)
    {
      double *p = stream + (((uint64_t)(it * KS + k) * nthreads + tid) & smask);
      *p = *p + 1.0;
    }
    // stores to words that are not in L2
#pragma unroll
    for (int k = 0; k < M; ++k) tbl[mx[k]] = acc - k;
  }
  if (acc == 123.456) *sink = acc;
}

template <int R, int H, int M, int KS>
static void run(const char *name, double *tbl, uint64_t tmask, double *stream, uint64_t smask, double *sink)
{
  const int threads = 65536, block = 64, iters = 50;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_mix<R, H, M, KS><<<threads / block, block>>>(tbl, tmask, stream, smask, 5, sink);     // warm-up
  cudaEventRecord(e0);
  k_mix<R, H, M, KS><<<threads / block, block>>>(tbl, tmask, stream, smask, iters, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double ns = ms * 1e6 / ((double)threads * iters);
  printf("%-58s R=%2d H=%2d M=%2d KS=%3d  %7.2f ms  %6.3f ns per thread-iteration  -> %.3e iterations/s  (%s)\n", name, R, H, M, KS, ms, ns,
         1e9 / ns, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
  const uint64_t tbytes = 16ull << 30, sbytes = 256ull << 20;
  double *tbl, *stream, *sink;
  if (cudaMalloc(&tbl, tbytes) != cudaSuccess || cudaMalloc(&stream, sbytes) != cudaSuccess || cudaMalloc(&sink, 8) != cudaSuccess)
  { printf("alloc failed\n"); return 1; }
  cudaMemset(tbl, 0, tbytes); cudaMemset(stream, 0, sbytes);
  const uint64_t tmask = tbytes / 8 - 1, smask = sbytes / 8 - 1;
  for (int rep = 0; rep < 2; ++rep)
  {
    run<64, 16, 14, 144>("today: 64 reads, 16 sentinel stores (L2 hits), 14 pops", tbl, tmask, stream, smask, sink);
    run<64, 0, 30, 144>("v8: the same with the sentinel stores after the sweep", tbl, tmask, stream, smask, sink);
    run<48, 10, 20, 144>("p registered at act time, sentinels of re-visits rewritten", tbl, tmask, stream, smask, sink);
    run<48, 10, 14, 144>("p registered at act time + stable slots", tbl, tmask, stream, smask, sink);
    run<38, 10, 14, 144>("compulsory fetches only", tbl, tmask, stream, smask, sink);
    run<64, 16, 14, 0>("today without the dense stream", tbl, tmask, stream, smask, sink);
    run<48, 10, 14, 0>("stable slots without the dense stream", tbl, tmask, stream, smask, sink);
    // marginal costs of the three kinds of operation
    run<64, 0, 14, 144>("today without the 16 L2-hit stores", tbl, tmask, stream, smask, sink);
    run<64, 16, 0, 144>("today without the 14 pops", tbl, tmask, stream, smask, sink);
    run<48, 16, 14, 144>("today without 16 of the reads", tbl, tmask, stream, smask, sink);
    // the actor-critic agent: 4 x 16 reads, 16 actor stores + 16 sentinels on words just read, ~12 pops, V = 64
    run<64, 32, 12, 64>("actor-critic today (k_run_ac)", tbl, tmask, stream, smask, sink);
    run<48, 26, 12, 64>("actor-critic, critic's p registered at act time + stable slots", tbl, tmask, stream, smask, sink);
  }
  return 0;
}
