// randbench5.cu -- would an L2-resident per-agent directory survive next to the kernel's DRAM traffic?
// Per thread and iteration: KD random 8-byte reads into a 16 GiB table (DRAM), KP random 4-byte probes into a
// directory region (64 MiB, optionally under a persisting L2 access-policy window), KS coalesced 8-byte
// read+write stream elements over a 256 MiB buffer (the dense trace-cache traffic).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o randbench5.bin randbench5.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

__device__ __forceinline__ uint64_t policy_first() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p; }
__device__ __forceinline__ uint64_t policy_last() { uint64_t p; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p; }

// the same mix with L2 eviction hints: table reads and the stream evict_first, the directory evict_last
template <int KD, int KP, int KS>
__global__ void k_mix_hint(const double *tbl, uint64_t tmask, uint32_t *dir, uint32_t dmask_per_thread, double *stream, uint64_t smask,
                           int iters, double *sink, int probe_writes)
{
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
  uint64_t s = (uint64_t)tid * 0x9E3779B97F4A7C15ull + 12345;
  double acc = 0;
  const uint64_t pf = policy_first(), pl = policy_last();
  uint32_t *mydir = dir + (size_t)tid * (dmask_per_thread + 1);
  for (int it = 0; it < iters; ++it)
  {
    uint64_t ix[KD > 0 ? KD : 1];
    uint32_t px[KP > 0 ? KP : 1];
#pragma unroll
    for (int k = 0; k < KD; ++k) { s = s * 6364136223846793005ull + 1442695040888963407ull; ix[k] = (s >> 17) & tmask; }
#pragma unroll
    for (int k = 0; k < KP; ++k) { s = s * 6364136223846793005ull + 1442695040888963407ull; px[k] = (uint32_t)(s >> 33) & dmask_per_thread; }
    double v[KD > 0 ? KD : 1];
    uint32_t w[KP > 0 ? KP : 1];
#pragma unroll
    for (int k = 0; k < KP; ++k) asm volatile("ld.global.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(w[k]) : "l"(mydir + px[k]), "l"(pl));
#pragma unroll
    for (int k = 0; k < KD; ++k) asm volatile("ld.global.L2::cache_hint.L2::64B.f64 %0, [%1], %2;" : "=d"(v[k]) : "l"(tbl + ix[k]), "l"(pf));
#pragma unroll
    for (int k = 0; k < KS; ++k)
    {
      double *p = stream + (((uint64_t)(it * KS + k) * nthreads + tid) & smask);
      double x;
      asm volatile("ld.global.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(x) : "l"(p), "l"(pf));
      x += 1.0;
      asm volatile("st.global.L2::cache_hint.f64 [%0], %1, %2;" :: "l"(p), "d"(x), "l"(pf) : "memory");
    }
#pragma unroll
    for (int k = 0; k < KD; ++k) acc += v[k];
#pragma unroll
    for (int k = 0; k < KP; ++k) acc += (double)w[k];
    if (probe_writes)
    {
#pragma unroll
      for (int k = 0; k < KP; k += 4) asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" :: "l"(mydir + px[k]), "r"((uint32_t)it), "l"(pl) : "memory");
    }
  }
  if (acc == 123.456) *sink = acc;
}

template <int KD, int KP, int KS>
__global__ void k_mix(const double *tbl, uint64_t tmask, uint32_t *dir, uint32_t dmask_per_thread, double *stream, uint64_t smask,
                      int iters, double *sink, int probe_writes)
{
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
  uint64_t s = (uint64_t)tid * 0x9E3779B97F4A7C15ull + 12345;
  double acc = 0;
  uint32_t *mydir = dir + (size_t)tid * (dmask_per_thread + 1);     // agent-major: 1 KiB per thread
  for (int it = 0; it < iters; ++it)
  {
    uint64_t ix[KD > 0 ? KD : 1];
    uint32_t px[KP > 0 ? KP : 1];
#pragma unroll
    for (int k = 0; k < KD; ++k) { s = s * 6364136223846793005ull + 1442695040888963407ull; ix[k] = (s >> 17) & tmask; }
#pragma unroll
    for (int k = 0; k < KP; ++k) { s = s * 6364136223846793005ull + 1442695040888963407ull; px[k] = (uint32_t)(s >> 33) & dmask_per_thread; }
    double v[KD > 0 ? KD : 1];
    uint32_t w[KP > 0 ? KP : 1];
#pragma unroll
    for (int k = 0; k < KP; ++k) w[k] = mydir[px[k]];
#pragma unroll
    for (int k = 0; k < KD; ++k) asm volatile("ld.global.L2::64B.f64 %0, [%1];" : "=d"(v[k]) : "l"(tbl + ix[k]));
#pragma unroll
    for (int k = 0; k < KS; ++k)
    {
      double *p = stream + (((uint64_t)(it * KS + k) * nthreads + tid) & smask);
      *p = *p + 1.0;
    }
#pragma unroll
    for (int k = 0; k < KD; ++k) acc += v[k];
#pragma unroll
    for (int k = 0; k < KP; ++k) acc += (double)w[k];
    if (probe_writes)
    {
#pragma unroll
      for (int k = 0; k < KP; k += 4) mydir[px[k]] = (uint32_t)it;
    }
  }
  if (acc == 123.456) *sink = acc;
}

static int g_hint = 0;
template <int KD, int KP, int KS>
static void run(const char *label, const double *tbl, uint64_t telems, uint32_t *dir, double *stream, uint64_t selems, double *sink,
                cudaStream_t st, int probe_writes)
{
  if (g_hint)
  {
    const int threads = 65536, iters = 64;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_mix_hint<KD, KP, KS><<<threads / 64, 64, 0, st>>>(tbl, telems - 1, dir, 255, stream, selems - 1, 4, sink, probe_writes);
    cudaEventRecord(e0, st);
    k_mix_hint<KD, KP, KS><<<threads / 64, 64, 0, st>>>(tbl, telems - 1, dir, 255, stream, selems - 1, iters, sink, probe_writes);
    cudaEventRecord(e1, st);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double per_iter_ns = ms * 1e6 / ((double)threads * iters);
    printf("%-34s KD=%2d KP=%2d KS=%3d w=%d  %7.2f ms  %6.3f ns per thread-iteration  DRAM-read rate %6.1f G/s  probe rate %6.1f G/s (%s)\n", label, KD, KP, KS,
           probe_writes, ms, per_iter_ns, KD / per_iter_ns, KP / per_iter_ns, cudaGetErrorString(cudaGetLastError()));
    return;
  }
  const int threads = 65536, iters = 64;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_mix<KD, KP, KS><<<threads / 64, 64, 0, st>>>(tbl, telems - 1, dir, 255, stream, selems - 1, 4, sink, probe_writes);
  cudaEventRecord(e0, st);
  k_mix<KD, KP, KS><<<threads / 64, 64, 0, st>>>(tbl, telems - 1, dir, 255, stream, selems - 1, iters, sink, probe_writes);
  cudaEventRecord(e1, st);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double per_iter_ns = ms * 1e6 / ((double)threads * iters);
  printf("%-34s KD=%2d KP=%2d KS=%3d w=%d  %7.2f ms  %6.3f ns per thread-iteration  DRAM-read rate %6.1f G/s  probe rate %6.1f G/s (%s)\n", label, KD, KP, KS,
         probe_writes, ms, per_iter_ns, KD / per_iter_ns, KP / per_iter_ns, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
  double *sink, *tbl, *stream;
  uint32_t *dir;
  cudaMalloc(&sink, 8);
  const uint64_t telems = 1ull << 31, selems = 1ull << 25;     // 16 GiB table, 256 MiB stream
  const size_t dir_bytes = (size_t)65536 * 256 * 4;           // 64 MiB directory: 1 KiB per thread
  if (cudaMalloc(&tbl, telems * 8) != cudaSuccess || cudaMalloc(&stream, selems * 8) != cudaSuccess ||
      cudaMalloc(&dir, dir_bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(tbl, 0, telems * 8); cudaMemset(stream, 0, selems * 8); cudaMemset(dir, 0, dir_bytes);
  int dev = 0, max_persist = 0, l2 = 0, max_window = 0;
  cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
  cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev);
  cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev);
  printf("L2 %d MiB, max persisting %d MiB, max access-policy window %d MiB\n", l2 >> 20, max_persist >> 20, max_window >> 20);
  cudaStream_t st;
  cudaStreamCreate(&st);
  for (int persist = 0; persist < 3; ++persist)
  {
    if (persist == 2)
    {
      cudaStreamAttrValue attr;
      attr.accessPolicyWindow.base_ptr = dir; attr.accessPolicyWindow.num_bytes = 0; attr.accessPolicyWindow.hitRatio = 0.f;
      attr.accessPolicyWindow.hitProp = cudaAccessPropertyNormal; attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
      cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &attr);
      cudaCtxResetPersistingL2Cache();
      g_hint = 1;
      printf("-- per-instruction L2 eviction hints: table reads + stream evict_first, directory evict_last\n");
    }
    else if (persist)
    {
      cudaError_t e = cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)max_persist);
      cudaStreamAttrValue attr;
      attr.accessPolicyWindow.base_ptr = dir;
      attr.accessPolicyWindow.num_bytes = dir_bytes < (size_t)max_window ? dir_bytes : (size_t)max_window;
      attr.accessPolicyWindow.hitRatio = 1.0f;
      attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
      attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      cudaError_t e2 = cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &attr);
      printf("-- persisting window on the directory: limit %s, attribute %s\n", cudaGetErrorString(e), cudaGetErrorString(e2));
    }
    else printf("-- no access policy\n");
    run<48, 0, 0>("DRAM reads only", tbl, telems, dir, stream, selems, sink, st, 0);
    run<0, 64, 0>("probes only", tbl, telems, dir, stream, selems, sink, st, 0);
    run<0, 64, 0>("probes only, 1/4 written", tbl, telems, dir, stream, selems, sink, st, 1);
    run<38, 64, 0>("DRAM reads + probes", tbl, telems, dir, stream, selems, sink, st, 0);
    run<38, 64, 128>("DRAM reads + probes + stream", tbl, telems, dir, stream, selems, sink, st, 0);
    run<38, 64, 128>("DRAM + probes(1/4 w) + stream", tbl, telems, dir, stream, selems, sink, st, 1);
    run<64, 0, 128>("today: 64 DRAM reads + stream", tbl, telems, dir, stream, selems, sink, st, 0);
    run<38, 0, 128>("38 DRAM reads + stream", tbl, telems, dir, stream, selems, sink, st, 0);
  }
  cudaDeviceSynchronize();
  return 0;
}
