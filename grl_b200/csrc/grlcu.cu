// grlcu.cu -- C ABI (include/grlcu.h) over the sm_100a kernels.  No CPU fallback: every entry point
// needs a usable CUDA device and reports GRLCU_ERR_CUDA otherwise.
#include "../../include/grlcu.h"
#include "grlcu_kernels.cuh"
#include "grlcu_ac.cuh"
#include "grlcu_layout.h"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>          // types only: the library is loaded with dlopen on first use (grlcu_group_*)
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdarg>
#include <cstring>
#include <new>
#include <vector>
#include <algorithm>
#include <numeric>

using namespace grlcu;

static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess)                                                                \
      return fail(GRLCU_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                  __FILE__, __LINE__);                                                    \
  } while (0)

struct grlcu_ctx
{
  grlcu_config cfg;
  Params p;
  ACParams ac;
  int device;
  cudaStream_t stream;
  int S, OBS;
  std::vector<void *> allocs;
  double *d_alpha, *d_lambda, *d_epsilon, *d_start;
  bool warned_slow;
  uint32_t *d_chunk_map;      // launch order of the 32-agent chunks (per-agent lambda), see grlcu_set_hyper
  int64_t launches;
  int sm_count;
  bool geo;                   // tile-major physical table layout in use (struct Geo)
  uint32_t inbox_words, shared_words;
  double *d_stage;            // [memory] staging row for logical <-> physical table copies
  bool cache_live;            // the trace value cache of the fused kernel is live (tables hold directory sentinels)
  // run-kernel timing (events recorded around every k_run launch)
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_pending;
  std::vector<cudaEvent_t> ev_pool;
  double ev_ms; int64_t ev_n;  // time / count of the launches already folded out of ev_pending
  bool traces_live;           // agents have stepped since the last GRLCU_RESET_AGENTS (traces may be non-empty)
  int block;
};

extern "C" const char *grlcu_last_error(void) { return g_err; }
extern "C" int grlcu_version(void) { return 1; }

// ---- random-access probe (include/grlcu.h): what this device does for the path's access pattern ---------------
template <int MODE>   // 0 read, 1 write, 2 read-modify-write
__global__ void k_probe(double *buf, uint64_t mask, int iters, double *sink)
{
  constexpr int K = 16;
  uint64_t s = (uint64_t)(blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345;
  double acc = 0;
  for (int it = 0; it < iters; ++it)
  {
    uint64_t ix[K];
    double v[K];
#pragma unroll
    for (int k = 0; k < K; ++k) { s = s * 6364136223846793005ull + 1442695040888963407ull; ix[k] = (s >> 17) & mask; }
    if (MODE != 1)
    {
#pragma unroll
      for (int k = 0; k < K; ++k) v[k] = ld_tbl(buf + ix[k]);
    }
    if (MODE == 0)
    {
#pragma unroll
      for (int k = 0; k < K; ++k) acc += v[k];
    }
    else
    {
#pragma unroll
      for (int k = 0; k < K; ++k) buf[ix[k]] = MODE == 1 ? (double)it : v[k] + 1.0;
    }
  }
  if (acc == 123.456) *sink = acc;
}

extern "C" int grlcu_probe_random_access(int32_t device, uint64_t footprint_bytes, double *reads_per_s, double *writes_per_s,
                                         double *rmws_per_s)
{
  if (footprint_bytes < (64ull << 20) || (footprint_bytes & (footprint_bytes - 1)))
    return fail(GRLCU_ERR_ARG, "grlcu_probe_random_access: footprint must be a power of two >= 64 MiB");
  CU(cudaSetDevice(device));
  double *buf = nullptr, *sink = nullptr;
  if (cudaMalloc(&buf, footprint_bytes) != cudaSuccess) { cudaGetLastError(); return fail(GRLCU_ERR_NOMEM, "grlcu_probe_random_access: no room for the scratch buffer"); }
  if (cudaMalloc(&sink, 8) != cudaSuccess) { cudaFree(buf); cudaGetLastError(); return fail(GRLCU_ERR_NOMEM, "grlcu_probe_random_access: out of device memory"); }
  cudaMemset(buf, 0, footprint_bytes);
  const uint64_t mask = footprint_bytes / 8 - 1;
  const int threads = 65536, iters = 64;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double *outs[3] = {reads_per_s, writes_per_s, rmws_per_s};
  for (int mode = 0; mode < 3; ++mode)
  {
    if (!outs[mode]) continue;
    for (int rep = 0; rep < 2; ++rep)      // first pass warms up
    {
      cudaEventRecord(e0);
      if (mode == 0) k_probe<0><<<threads / 64, 64>>>(buf, mask, rep ? iters : 2, sink);
      else if (mode == 1) k_probe<1><<<threads / 64, 64>>>(buf, mask, rep ? iters : 2, sink);
      else k_probe<2><<<threads / 64, 64>>>(buf, mask, rep ? iters : 2, sink);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    *outs[mode] = ms > 0 ? (double)threads * iters * 16 / (ms * 1e-3) : 0.;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(buf); cudaFree(sink);
  CU(cudaGetLastError());
  return 0;
}

template <typename Tp>
static int dalloc(grlcu_ctx *c, Tp **ptr, size_t count)
{
  void *q = nullptr;
  cudaError_t e = cudaMalloc(&q, count * sizeof(Tp));
  if (e != cudaSuccess)
    return fail(GRLCU_ERR_NOMEM, "cudaMalloc of %zu bytes failed: %s", count * sizeof(Tp), cudaGetErrorString(e));
  c->allocs.push_back(q);
  *ptr = (Tp *)q;
  return 0;
}

static int grid_for(int64_t work, int block, int cap = 148 * 16)
{
  int64_t g = (work + block - 1) / block;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

static int supported(const grlcu_config *cfg)
{
  if (cfg->tilings != 16) return fail(GRLCU_ERR_UNSUPPORTED, "tilings %d: kernels are instantiated for 16", cfg->tilings);
  if (cfg->memory < 1) return fail(GRLCU_ERR_ARG, "memory must be positive");
  if (cfg->integration_steps < 1) return fail(GRLCU_ERR_ARG, "integration_steps must be >= 1");
  if (cfg->end_stop_penalty != 0 && cfg->end_stop_penalty != 1) return fail(GRLCU_ERR_ARG, "end_stop_penalty must be 0 or 1");
  if (cfg->safe < 0 || cfg->safe > 2) return fail(GRLCU_ERR_ARG, "projector/tile_coding:safe must be 0, 1 or 2");
  if (cfg->safe > 0 && cfg->algo != GRLCU_ALGO_SARSA)
    return fail(GRLCU_ERR_UNSUPPORTED, "safe > 0 (the collision-claim table) is instantiated for predictor/sarsa only");
  if (cfg->algo == GRLCU_ALGO_AC)
  {
    // cfg/cart_double_pole/ac.yaml, cfg/pendulum/ac_tc.yaml, cfg/cart_pole/ac_tc.yaml of the reference
    if (cfg->dynamics != GRLCU_DYN_PENDULUM && cfg->dynamics != GRLCU_DYN_CART_POLE && cfg->dynamics != GRLCU_DYN_CART_DOUBLE_POLE)
      return fail(GRLCU_ERR_UNSUPPORTED, "dynamics %d not instantiated for actor-critic", cfg->dynamics);
    const int aobs = cfg->dynamics == GRLCU_DYN_PENDULUM ? 2 : cfg->dynamics == GRLCU_DYN_CART_POLE ? 4 : 6;
    if (cfg->obs_dims != aobs) return fail(GRLCU_ERR_ARG, "obs_dims %d does not match dynamics (%d)", cfg->obs_dims, aobs);
    for (int i = 0; i < cfg->obs_dims; ++i)
      if (cfg->actor_resolution[i] != cfg->resolution[i] || cfg->actor_wrapping[i] != cfg->wrapping[i])
        return fail(GRLCU_ERR_UNSUPPORTED, "actor and critic projectors must have the same resolution and wrapping "
                    "(one index set addresses both tables; cfg/cart_double_pole/ac.yaml)");
    if (!(cfg->action_min < cfg->action_max)) return fail(GRLCU_ERR_ARG, "action_min must be below action_max");
    if (cfg->sigma < 0) return fail(GRLCU_ERR_ARG, "sigma must be >= 0");
    return 0;
  }
  if (cfg->algo != GRLCU_ALGO_SARSA && cfg->algo != GRLCU_ALGO_Q && cfg->algo != GRLCU_ALGO_EXPECTED_SARSA)
    return fail(GRLCU_ERR_UNSUPPORTED, "unknown algo %d", cfg->algo);
  if (cfg->algo == GRLCU_ALGO_EXPECTED_SARSA && cfg->target_epsilon > 1) return fail(GRLCU_ERR_ARG, "target_epsilon > 1");
  if (cfg->sampler != GRLCU_SAMPLER_EPSILON_GREEDY && cfg->sampler != GRLCU_SAMPLER_SOFTMAX)
    return fail(GRLCU_ERR_UNSUPPORTED, "unknown sampler %d", cfg->sampler);
  // softmax.cpp:34: tau in [1e-6, 100]
  if (cfg->sampler == GRLCU_SAMPLER_SOFTMAX && !(cfg->softmax_tau >= 0.000001 && cfg->softmax_tau <= 100.))
    return fail(GRLCU_ERR_ARG, "softmax_tau must be in [1e-6, 100]");
  // cfg/pendulum/sarsa_tc.yaml, cfg/cart_pole/sarsa.yaml, cfg/cart_double_pole/sarsa.yaml of the reference
  if (cfg->dynamics != GRLCU_DYN_PENDULUM && cfg->dynamics != GRLCU_DYN_CART_POLE && cfg->dynamics != GRLCU_DYN_CART_DOUBLE_POLE)
    return fail(GRLCU_ERR_UNSUPPORTED, "dynamics %d not instantiated for SARSA", cfg->dynamics);
  // discretizer/uniform with one action dimension (uniform.cpp:126-151): 3 actions run every kernel (value cache, tile-major
  // layout); 2, 4 and 5 are instantiated for the uncached kernels only
  if (cfg->n_actions < 2 || cfg->n_actions > 5)
    return fail(GRLCU_ERR_UNSUPPORTED, "n_actions %d: kernels are instantiated for 2..5 actions (one action dimension)", cfg->n_actions);
  if (cfg->n_actions != 3 && cfg->safe > 0) return fail(GRLCU_ERR_UNSUPPORTED, "safe > 0 is instantiated for 3 actions");
  const int obs = cfg->dynamics == GRLCU_DYN_PENDULUM ? 2 : cfg->dynamics == GRLCU_DYN_CART_POLE ? 4 : 6;
  if (cfg->obs_dims != obs) return fail(GRLCU_ERR_ARG, "obs_dims %d does not match dynamics (%d)", cfg->obs_dims, obs);
  return 0;
}

// smallest n with e^n < dc by repeated multiplication (the trace's own arithmetic) + 1 slack
static int trace_capacity_for(double e, int kind)
{
  if (kind == GRLCU_TRACE_NONE) return 1;
  const double dc = kind == GRLCU_TRACE_REPLACING ? 0.01 : 0.0001;
  if (!(e < 1.0)) return -1;
  double t = 1.;
  int n = 0;
  while (!(t < dc) && n < 4096) { t *= e; ++n; }
  return n + 1;
}

static int launch_reset(grlcu_ctx *c, int what);
static int ensure_plain(grlcu_ctx *c);

// ---- pilot run for the visit-ordered table layout (grlcu_layout.h: build_layout_from_visits) --------------------------
// Which groups of tiles the agents of this configuration actually visit is found by running a few of them: a small
// context in the raw layout whose projections are logged (Geo::pilot_keys).  All agents of a context are statistically
// alike (same task, same hyper-parameters), so a few hundred predict the rest; a group the pilot never saw merely is not
// co-located.  The pilot's agents are thrown away; the real agents start from scratch.
static thread_local int g_in_pilot = 0;

static int run_pilot(const grlcu_config *cfg, int64_t first_agent_id, int device, void *stream, std::vector<uint64_t> *keys)
{
  const size_t budget = (size_t)2 << 30;                          // device memory the pilot's tables may take
  int agents = (int)std::min<size_t>(256, budget / ((size_t)cfg->memory * sizeof(double)));
  if (agents < 32) return 0;
  const int steps = getenv("GRLCU_PILOT_STEPS") ? std::max(100, atoi(getenv("GRLCU_PILOT_STEPS"))) : 600;
  const uint32_t cap = (uint32_t)(steps + 16) * (uint32_t)cfg->tilings;      // keys per agent
  grlcu_ctx *pc = nullptr;
  g_in_pilot = 1;
  int r = grlcu_create(cfg, agents, first_agent_id, device, stream, &pc);
  g_in_pilot = 0;
  if (r) return r;
  uint64_t *d_keys = nullptr;
  uint32_t *d_count = nullptr;
  r = dalloc(pc, &d_keys, (size_t)cap * agents);
  if (!r) r = dalloc(pc, &d_count, (size_t)agents);
  if (!r && cudaMemsetAsync(d_count, 0, sizeof(uint32_t) * agents, pc->stream) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "cudaMemset failed");
  if (!r)
  {
    pc->p.geo.pilot_keys = d_keys; pc->p.geo.pilot_count = d_count; pc->p.geo.pilot_cap = cap;
    r = grlcu_run(pc, steps, 0);
  }
  std::vector<uint32_t> counts(agents, 0);
  std::vector<uint64_t> all;
  if (!r && cudaStreamSynchronize(pc->stream) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "pilot run failed: %s", cudaGetErrorString(cudaGetLastError()));
  if (!r && cudaMemcpy(counts.data(), d_count, sizeof(uint32_t) * agents, cudaMemcpyDeviceToHost) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "pilot: copy failed");
  if (!r)
  {
    all.resize((size_t)cap * agents);
    if (cudaMemcpy(all.data(), d_keys, all.size() * sizeof(uint64_t), cudaMemcpyDeviceToHost) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "pilot: copy failed");
  }
  if (!r)
  {
    // agent after agent, each in its own projection order
    keys->clear();
    for (int a = 0; a < agents; ++a)
      keys->insert(keys->end(), all.begin() + (size_t)a * cap, all.begin() + (size_t)a * cap + std::min(counts[a], cap));
  }
  grlcu_destroy(pc);
  return r;
}

extern "C" int grlcu_create(const grlcu_config *cfg, int32_t n_agents, int64_t first_agent_id,
                            int32_t device, void *stream, grlcu_ctx **out)
{
  if (!cfg || !out || n_agents < 1) return fail(GRLCU_ERR_ARG, "grlcu_create: bad arguments");
  *out = nullptr;
  if (int r = supported(cfg)) return r;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev < 1)
    return fail(GRLCU_ERR_CUDA, "no CUDA device available (%s); libgrlcu has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail(GRLCU_ERR_ARG, "device %d out of range (%d devices)", device, ndev);
  CU(cudaSetDevice(device));
  // (cudaLimitMaxL2FetchGranularity was measured to change nothing for the hashed gathers and is left alone: a
  // library has no business changing device-wide limits; ld_tbl() asks for 64-byte fetches per instruction instead)

  grlcu_ctx *c = new (std::nothrow) grlcu_ctx();
  if (!c) return fail(GRLCU_ERR_NOMEM, "out of host memory");
  c->cfg = *cfg;
  c->device = device;
  c->stream = (cudaStream_t)stream;
  c->S = cfg->dynamics == GRLCU_DYN_PENDULUM ? 3 : cfg->dynamics == GRLCU_DYN_CART_POLE ? 5 : 7;
  c->OBS = cfg->obs_dims;
  c->d_alpha = c->d_lambda = c->d_epsilon = c->d_start = nullptr;
  c->d_chunk_map = nullptr; c->warned_slow = false;
  c->launches = 0;
  c->ev_ms = 0; c->ev_n = 0; c->traces_live = false;
  c->geo = false; c->inbox_words = c->shared_words = 0; c->d_stage = nullptr;
  c->cache_live = false;
  c->block = 64;
  c->sm_count = 148;
  cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
  if (const char *b = getenv("GRLCU_BLOCK")) { int v = atoi(b); if (v >= 32 && v <= kMaxBlock && v % 32 == 0) c->block = v; }

  Params &p = c->p;
  memset(&p, 0, sizeof(p));
  p.n = n_agents; p.first_id = first_agent_id; p.seed = cfg->seed;
  p.tau = cfg->control_step; p.isteps = cfg->integration_steps;
  p.timeout = cfg->timeout; p.randomization = cfg->randomization; p.end_stop_penalty = cfg->end_stop_penalty;
  // TileCodingProjector::configure (tile_coding.cpp:66-79)
  const bool is_ac = cfg->algo == GRLCU_ALGO_AC;
  const int dims = is_ac ? cfg->obs_dims : cfg->obs_dims + 1;   // AC projects observations, SARSA (obs, action)
  for (int i = 0; i < dims; ++i)
  {
    p.proj.scaling[i] = cfg->tilings / cfg->resolution[i];
    double w = cfg->wrapping[i] * p.proj.scaling[i];
    if (fabs(w - round(w)) > 0.001)
    {
      delete c;
      return fail(GRLCU_ERR_ARG, "projector/tile_coding:wrapping: scaled wrapping for dimension %d (%.5f) is not an integer", i, w);
    }
    p.proj.wrap[i] = (int32_t)round(w);
  }
  p.proj.memory = cfg->memory;
  // Words of the tile hash that do not depend on the state, pre-mixed exactly as createHashSum does
  // (tile_coding.h:96-106): the action coordinate of tiling j (tile_coding.cpp:122-140) and j itself.
  {
    const uint32_t m = 0x5bd1e995u;
    auto mix = [m](uint32_t k) { k *= m; k ^= k >> 24; k *= m; return k; };
    const int ia = cfg->obs_dims;                 // the action is the last projector input
    for (int aa = 0; aa < (is_ac ? 0 : cfg->n_actions); ++aa)
    {
      const int32_t qa = (int32_t)floor(cfg->actions[aa] * p.proj.scaling[ia]);
      for (int j = 0; j < cfg->tilings; ++j)
      {
        const int32_t base = j * (1 + 2 * ia);
        int32_t r = (qa - base) % cfg->tilings; if (r < 0) r += cfg->tilings;
        int32_t cc = qa - r;
        if (p.proj.wrap[ia] != 0) { cc %= p.proj.wrap[ia]; if (cc < 0) cc += p.proj.wrap[ia]; }
        p.kact[aa][j] = mix((uint32_t)cc);
      }
    }
    for (int j = 0; j < cfg->tilings; ++j) p.kj[j] = mix((uint32_t)j);
    const uint32_t mem = (uint32_t)cfg->memory;
    p.mem_mask = (mem & (mem - 1)) == 0 ? mem - 1 : 0;
    if (mem == 1) p.mem_mask = 0;
    p.mem_magic = mem > 1 ? 0xFFFFFFFFFFFFFFFFull / mem + 1 : 0;
  }
  for (int i = 0; i < cfg->n_actions; ++i) p.actions[i] = cfg->actions[i];
  p.test_interval = cfg->test_interval;
  p.alpha = cfg->alpha; p.gamma = cfg->gamma; p.lambda = cfg->lambda; p.epsilon = cfg->epsilon;
  p.trace_kind = cfg->trace;
  p.force_slow = getenv("GRLCU_FORCE_SLOW") ? atoi(getenv("GRLCU_FORCE_SLOW")) : 0;
  p.mode = getenv("GRLCU_MODE") ? atoi(getenv("GRLCU_MODE")) : 0;
  p.target_kind = cfg->algo == GRLCU_ALGO_Q ? 1 : cfg->algo == GRLCU_ALGO_EXPECTED_SARSA ? 2 : 0;
  p.target_epsilon = cfg->target_epsilon;
  p.softmax_tau = cfg->sampler == GRLCU_SAMPLER_SOFTMAX ? cfg->softmax_tau : 0.;
  p.safe = cfg->safe;
  if (p.safe > 0) { p.mode = 1; p.force_slow = 1; }     // claim table: the reference's literal sequential update, uncached
  p.trace_cap = trace_capacity_for(cfg->gamma * cfg->lambda, cfg->trace);
  if (p.trace_cap < 0 || p.trace_cap > 160)
  {
    delete c;
    return fail(GRLCU_ERR_UNSUPPORTED, "gamma*lambda = %g needs an unbounded or too long trace", cfg->gamma * cfg->lambda);
  }

  // physical table layout: tile-major where the task's tiles fit the table (GRLCU_LAYOUT=0: the reference's raw
  // hash order, kept for A/B runs and as the other side of tests/test_gpu_cached.py)
  p.tstride = (uint32_t)cfg->memory;
  grlcu_host::Layout lay;
  const int layout_env = getenv("GRLCU_LAYOUT") ? atoi(getenv("GRLCU_LAYOUT")) : 1;     // 0 raw, 2 visit-ordered even where tile-major applies
  if (layout_env == 1 && cfg->safe == 0 && !g_in_pilot)      // (the pilot runs in the raw layout)
    lay = grlcu_host::build_layout(*cfg, p.proj.scaling, p.proj.wrap);
  const bool pilot_ok = !g_in_pilot && layout_env != 0 && cfg->safe == 0 && !is_ac && cfg->n_actions == 3 && cfg->tilings == 16 &&
                        (cfg->memory & 3) == 0 && cfg->memory >= 4096 && cfg->memory <= (1 << 20) && n_agents >= 64;
  // a box that fills more than half the table leaves many tiles without a home (sweep: 30 % of the words share a slot):
  // ranking by actual visits does better there (4.47e8 against 4.22e8), profiles/r02_visit_ordered_vs_tile_major.txt
  if (lay.enabled && pilot_ok && (size_t)lay.inbox_words * 2 > (size_t)cfg->memory) lay = grlcu_host::Layout();
  if (lay.enabled)
  {
    c->geo = true; c->inbox_words = lay.inbox_words; c->shared_words = lay.shared_words;
    p.tstride = lay.tstride;
    p.geo.enabled = 1;
    for (int i = 0; i < cfg->obs_dims; ++i) { p.geo.kn[i] = lay.kn[i]; p.geo.koff[i] = lay.koff[i]; p.geo.wrapped[i] = lay.wrapped[i]; }
  }
  else if (pilot_ok)
  {
    // the task's tiles do not fit the table (or it has no box): visit-ordered layout from a pilot run.  A failing pilot
    // only costs the layout.
    std::vector<uint64_t> keys;
    if (run_pilot(cfg, first_agent_id, device, stream, &keys) == 0)
      lay = grlcu_host::build_layout_from_visits(*cfg, keys, p.kact, p.kj);
    CU(cudaSetDevice(device));
    if (lay.enabled)
    {
      c->geo = true; c->inbox_words = lay.inbox_words; c->shared_words = lay.shared_words;
      p.tstride = lay.tstride;
      p.geo.enabled = 2;
    }
    else if (layout_env == 1 && (lay = grlcu_host::build_layout(*cfg, p.proj.scaling, p.proj.wrap)).enabled)
    {
      // (the pilot failed: tile-major after all, where it applies)
      c->geo = true; c->inbox_words = lay.inbox_words; c->shared_words = lay.shared_words;
      p.tstride = lay.tstride;
      p.geo.enabled = 1;
      for (int i = 0; i < cfg->obs_dims; ++i) { p.geo.kn[i] = lay.kn[i]; p.geo.koff[i] = lay.koff[i]; p.geo.wrapped[i] = lay.wrapped[i]; }
    }
  }

  const size_t n = (size_t)n_agents, T = (size_t)cfg->tilings;
  if (n * T * (size_t)p.trace_cap >= (1ull << 31))
  {
    delete c;
    return fail(GRLCU_ERR_UNSUPPORTED, "n_agents * tilings * trace capacity (%d) must stay below 2^31 per context", p.trace_cap);
  }
  int r = 0;
  if (!r && lay.enabled)
  {
    auto up = [&](const uint32_t **dst, const std::vector<uint32_t> &v) -> int
    {
      uint32_t *d = nullptr;
      if (int e = dalloc(c, &d, v.size())) return e;
      if (cudaMemcpy(d, v.data(), v.size() * sizeof(uint32_t), cudaMemcpyHostToDevice) != cudaSuccess)
        return fail(GRLCU_ERR_CUDA, "uploading the table layout failed");
      *dst = d;
      return 0;
    };
    if (!r) r = up(&p.geo.perm, lay.perm);
    if (!r) r = up(&p.geo.iperm, lay.iperm);
    if (!r && p.geo.enabled == 1) r = up(&p.geo.slotmap, lay.slotmap);
    if (!r && p.geo.enabled == 1) r = up(&p.geo.gflags, lay.gflags);
    if (!r) r = dalloc(c, &p.geo.mismatch, 1);
    if (!r && cudaMemset(p.geo.mismatch, 0, sizeof(int32_t)) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "cudaMemset failed");
    if (!r) r = dalloc(c, &c->d_stage, (size_t)cfg->memory);
  }
  if (is_ac) p.tstride = 2u * (uint32_t)cfg->memory;      // interleaved {critic, actor} pairs (grlcu_ac.cuh)
  if (!r) r = dalloc(c, &p.table, n * (size_t)p.tstride);
  memset(&c->ac, 0, sizeof(c->ac));
  if (!r && is_ac)
  {
    c->ac.actor = p.table + 1;
    r = dalloc(c, &c->d_stage, (size_t)cfg->memory);
    c->ac.sigma = cfg->sigma; c->ac.actor_alpha = cfg->actor_alpha;
    c->ac.action_min = cfg->action_min; c->ac.action_max = cfg->action_max;
  }
  if (!r && p.safe > 0)
  {
    r = dalloc(c, &p.claims, n * (size_t)cfg->memory);
    if (!r) r = dalloc(c, &p.prev_hash, n * T);
    if (!r) r = dalloc(c, &p.pressure, n);
    if (!r && cudaMemset(p.pressure, 0, n * sizeof(uint32_t)) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "cudaMemset failed");
    if (!r && cudaMemset(p.prev_hash, 0, n * T * sizeof(uint32_t)) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "cudaMemset failed");
  }
  if (!r) r = dalloc(c, &p.oob, 1);
  if (!r && cudaMemset(p.oob, 0, sizeof(int32_t)) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "cudaMemset failed");
  if (!r) r = dalloc(c, &p.state, n * c->S);
  if (!r) r = dalloc(c, &p.action, n);
  if (!r) r = dalloc(c, &p.prev_idx, n * T);
  if (!r && cudaMemset(p.prev_idx, 0, n * T * sizeof(uint32_t)) != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "cudaMemset failed");
  if (!r) r = dalloc(c, &p.trace_idx, warp_rows_agents(n) * T * (size_t)p.trace_cap);     // blocked by warp (warp_rows)
  if (!r) r = dalloc(c, &p.trace_len, n);
  if (!r) r = dalloc(c, &p.trace_head, n);
  if (!r) r = dalloc(c, &p.total_decay, n);
  if (!r) r = dalloc(c, &p.flags, n);
  if (!r) r = dalloc(c, &p.dupmask, n);
  {
    // cblock: agents per launch cohort (see grlcu_run); cache_val itself is blocked by warp (grlcu_cached.cuh)
    const int roomy_cap = c->sm_count * 8 * 32 / c->block * c->block;
    p.cblock = (uint32_t)n_agents;
    if (n_agents > roomy_cap)
    {
      const int parts = (n_agents + roomy_cap - 1) / roomy_cap;
      p.cblock = (uint32_t)(((n_agents + parts - 1) / parts + c->block - 1) / c->block * c->block);
    }
    if (!r) r = dalloc(c, &p.cache_val, warp_rows_agents(n) * T * (size_t)p.trace_cap);      // [n / 32][cap*T][32]
  }
  if (!r) r = dalloc(c, &p.cache_mask, warp_rows_agents(n) * (size_t)p.trace_cap);
  if (!r) r = dalloc(c, &p.act_ctr, n);
  if (!r) r = dalloc(c, &p.test_act_ctr, n);
  if (!r) r = dalloc(c, &p.episode_ctr, n);
  if (!r) r = dalloc(c, &p.tt, n);
  if (!r) r = dalloc(c, &p.ss, n);
  if (!r) r = dalloc(c, &p.swept, n);
  if (!r) r = dalloc(c, &p.tblw, n);
  if (!r) r = dalloc(c, &p.episodes, n);
  if (!r) r = dalloc(c, &p.total_reward, n);
  if (!r) r = dalloc(c, &p.last_return, n);
  if (!r) r = dalloc(c, &p.last_test_return, n);
  if (!r) r = dalloc(c, &p.return_sum, n);
  if (!r) r = launch_reset(c, GRLCU_RESET_TABLES | GRLCU_RESET_AGENTS | GRLCU_RESET_STATS);
  if (!r)
  {
    cudaError_t es = cudaStreamSynchronize(c->stream);
    if (es != cudaSuccess) r = fail(GRLCU_ERR_CUDA, "table initialisation failed: %s", cudaGetErrorString(es));
  }
  if (r) { grlcu_destroy(c); return r; }
  *out = c;
  return 0;
}

extern "C" int grlcu_destroy(grlcu_ctx *c)
{
  if (!c) return 0;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  for (auto &pr : c->ev_pending) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  for (auto ev : c->ev_pool) cudaEventDestroy(ev);
  for (void *q : c->allocs) cudaFree(q);
  delete c;
  return 0;
}

// ---- reset ---------------------------------------------------------------------------------
__global__ void k_reset_agents(Params p)
{
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= p.n) return;
  p.flags[a] = 0; p.dupmask[a] = 0; p.trace_len[a] = 0; p.trace_head[a] = 0; p.total_decay[a] = 1.;
  p.total_reward[a] = 0; p.action[a] = 0;
  p.act_ctr[a] = 0; p.test_act_ctr[a] = 0; p.episode_ctr[a] = 0; p.tt[a] = 0;
}

__global__ void k_reset_stats(Params p)
{
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= p.n) return;
  p.ss[a] = 0; p.swept[a] = 0; p.tblw[a] = 0; p.episodes[a] = 0;
  p.last_return[a] = 0; p.last_test_return[a] = 0; p.return_sum[a] = 0;
}

__global__ void k_begin_override(Params p)
{
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= p.n) return;
  p.flags[a] |= 4;     // start-state override pending: consumed by the next episode this agent begins (k_run)
}

static int launch_reset(grlcu_ctx *c, int what)
{
  Params &p = c->p;
  const int blk = 256, grd = (p.n + blk - 1) / blk;
  if (what & GRLCU_RESET_TABLES)
  {
    const int64_t work = ((int64_t)c->cfg.memory + 1) / 2 * p.n;
    if (c->ac.actor)
      k_table_init_pairs<<<grid_for((int64_t)c->cfg.memory * p.n, 256, 148 * 32), 256, 0, c->stream>>>(
          p.table, p.n, c->cfg.memory, p.first_id, p.seed, c->cfg.init_min, c->cfg.init_max);
    else if (c->geo)
      k_table_init_phys<<<grid_for((int64_t)p.tstride * p.n, 256, 148 * 32), 256, 0, c->stream>>>(
          p.table, p.n, p.tstride, p.geo.iperm, p.first_id, p.seed, c->cfg.init_min, c->cfg.init_max, STREAM_TABLE);
    else
      k_table_init<<<grid_for(work, 256, 148 * 32), 256, 0, c->stream>>>(
          p.table, p.n, c->cfg.memory, p.first_id, p.seed, c->cfg.init_min, c->cfg.init_max, STREAM_TABLE);
    c->launches++;
  }
  if ((what & GRLCU_RESET_TABLES) && p.claims)       // TileCodingProjector::reconfigure(action=reset), tile_coding.cpp:82-89
    CU(cudaMemsetAsync(p.claims, 0xFF, (size_t)p.n * c->cfg.memory * sizeof(int32_t), c->stream));
  if (what & GRLCU_RESET_AGENTS) { k_reset_agents<<<grd, blk, 0, c->stream>>>(p); c->launches++; c->traces_live = false; }
  if (what & GRLCU_RESET_STATS) { k_reset_stats<<<grd, blk, 0, c->stream>>>(p); c->launches++; }
  CU(cudaGetLastError());
  return 0;
}

// Tables and trace rings in the plain representation the reference (and every other entry point) expects.
static int ensure_plain(grlcu_ctx *c)
{
  if (!c->cache_live) return 0;
  const int blk = 64, grd = (c->p.n + blk - 1) / blk;
  if (c->cfg.dynamics == GRLCU_DYN_PENDULUM) k_cache_close<0, 16><<<grd, blk, 0, c->stream>>>(c->p);
  else if (c->cfg.dynamics == GRLCU_DYN_CART_POLE) k_cache_close<1, 16><<<grd, blk, 0, c->stream>>>(c->p);
  else k_cache_close<2, 16><<<grd, blk, 0, c->stream>>>(c->p);
  c->launches++;
  CU(cudaGetLastError());
  c->cache_live = false;
  return 0;
}

extern "C" int grlcu_reset(grlcu_ctx *c, int32_t what)
{
  if (!c) return fail(GRLCU_ERR_ARG, "null context");
  CU(cudaSetDevice(c->device));
  if (int r = ensure_plain(c)) return r;
  return launch_reset(c, what);
}

static int upload_opt(grlcu_ctx *c, double **dst, const double *src)
{
  if (!src) return 0;
  if (!*dst) if (int r = dalloc(c, dst, (size_t)c->p.n)) return r;
  CU(cudaMemcpyAsync(*dst, src, sizeof(double) * c->p.n, cudaMemcpyHostToDevice, c->stream));
  return 0;
}

extern "C" int grlcu_set_hyper(grlcu_ctx *c, const double *alpha, const double *lambda, const double *epsilon)
{
  if (!c) return fail(GRLCU_ERR_ARG, "null context");
  CU(cudaSetDevice(c->device));
  if (lambda)
  {
    // the trace ring stores no per-entry decay: the sweep and the pops would apply the NEW gamma*lambda to entries
    // pushed under the old one (the reference divides by front().decay, trace.h:229-233)
    if (c->traces_live)
      return fail(GRLCU_ERR_ARG, "grlcu_set_hyper: per-agent lambda can only change while every trace is empty "
                  "(right after grlcu_create or grlcu_reset(GRLCU_RESET_AGENTS))");
    double mx = 0;
    for (int i = 0; i < c->p.n; ++i) mx = lambda[i] > mx ? lambda[i] : mx;
    int cap = trace_capacity_for(c->cfg.gamma * mx, c->cfg.trace);
    if (cap < 0 || cap > c->p.trace_cap)
      return fail(GRLCU_ERR_ARG, "per-agent lambda up to %g needs trace capacity %d > %d: create the context "
                  "with config lambda = the largest lambda of the sweep", mx, cap, c->p.trace_cap);
  }
  if (int r = upload_opt(c, &c->d_alpha, alpha)) return r;
  if (int r = upload_opt(c, &c->d_lambda, lambda)) return r;
  if (int r = upload_opt(c, &c->d_epsilon, epsilon)) return r;
  if (lambda && c->cfg.algo != GRLCU_ALGO_AC && c->p.n % 32 == 0 && !getenv("GRLCU_NO_CHUNK_MAP"))
  {
    // A warp sweeps as long as its longest trace, and a cohort launch lasts as long as its slowest warp: order the 32-agent
    // chunks by (their largest) lambda so that the cohorts of grlcu_run hold traces of similar length and can use the
    // 255-register instantiation (one launch of everything needs the 128-register one and spills).
    const int chunks = c->p.n / 32;
    std::vector<double> key(chunks, 0.);
    for (int i = 0; i < c->p.n; ++i) key[i / 32] = std::max(key[i / 32], lambda[i]);
    std::vector<uint32_t> order(chunks);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return key[a] < key[b]; });
    if (!c->d_chunk_map) if (int r = dalloc(c, &c->d_chunk_map, (size_t)chunks)) return r;
    CU(cudaMemcpyAsync(c->d_chunk_map, order.data(), sizeof(uint32_t) * chunks, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));       // (order is a local)
    c->p.chunk_map = c->d_chunk_map;
  }
  CU(cudaStreamSynchronize(c->stream));
  c->p.alpha_a = c->d_alpha; c->p.lambda_a = c->d_lambda; c->p.epsilon_a = c->d_epsilon;
  return 0;
}

// ---- scratch helpers for the host-buffer entry points ------------------------------------------
struct Scratch
{
  std::vector<void *> ptrs;
  ~Scratch() { for (void *q : ptrs) cudaFree(q); }
  template <typename Tp> Tp *get(size_t count)
  {
    void *q = nullptr;
    if (cudaMalloc(&q, count * sizeof(Tp) + 16) != cudaSuccess) return nullptr;
    ptrs.push_back(q);
    return (Tp *)q;
  }
};

#define DISPATCH_DYN(c, CALL)                                        \
  do {                                                               \
    if ((c)->cfg.dynamics == GRLCU_DYN_PENDULUM) { CALL(0); }        \
    else if ((c)->cfg.dynamics == GRLCU_DYN_CART_POLE) { CALL(1); }  \
    else { CALL(2); }                                                \
  } while (0)
#define DISPATCH_SARSA(c, CALL) DISPATCH_DYN(c, CALL)

extern "C" int grlcu_project(grlcu_ctx *c, int64_t n, const double *in, uint32_t *hash, uint32_t *index)
{
  if (!c || n < 0 || !in || !index) return fail(GRLCU_ERR_ARG, "grlcu_project: bad arguments");
  if (n == 0) return 0;
  CU(cudaSetDevice(c->device));
  const bool is_ac = c->cfg.algo == GRLCU_ALGO_AC;
  const int dims = is_ac ? c->cfg.obs_dims : c->cfg.obs_dims + 1, T = c->cfg.tilings;
  Scratch s;
  double *d_in = s.get<double>((size_t)n * dims);
  uint32_t *d_h = s.get<uint32_t>((size_t)n * T), *d_i = s.get<uint32_t>((size_t)n * T);
  if (!d_in || !d_h || !d_i) return fail(GRLCU_ERR_NOMEM, "grlcu_project: out of device memory");
  CU(cudaMemcpyAsync(d_in, in, sizeof(double) * n * dims, cudaMemcpyHostToDevice, c->stream));
  const int grd = grid_for(n, 128);
#define CALL(D) k_project<Dyn<D>::OBS + 1, 16><<<grd, 128, 0, c->stream>>>(c->p.proj, n, d_in, d_h, d_i)
#define CALL_AC(D) k_project<Dyn<D>::OBS, 16><<<grd, 128, 0, c->stream>>>(c->p.proj, n, d_in, d_h, d_i)
  if (is_ac) DISPATCH_DYN(c, CALL_AC);
  else DISPATCH_SARSA(c, CALL);
#undef CALL_AC
#undef CALL
  c->launches++;
  CU(cudaGetLastError());
  if (hash) CU(cudaMemcpyAsync(hash, d_h, sizeof(uint32_t) * n * T, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(index, d_i, sizeof(uint32_t) * n * T, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int grlcu_model_step(grlcu_ctx *c, int64_t n, const double *state, const double *action,
                                double *next, double *obs, double *reward, int32_t *terminal)
{
  if (!c || n < 0 || !state || !action || !next || !obs || !reward || !terminal)
    return fail(GRLCU_ERR_ARG, "grlcu_model_step: bad arguments");
  if (n == 0) return 0;
  CU(cudaSetDevice(c->device));
  const int S = c->S, OBS = c->OBS;
  Scratch s;
  double *d_s = s.get<double>((size_t)n * S), *d_a = s.get<double>(n), *d_n = s.get<double>((size_t)n * S),
         *d_o = s.get<double>((size_t)n * OBS), *d_r = s.get<double>(n);
  int32_t *d_t = s.get<int32_t>(n);
  if (!d_s || !d_a || !d_n || !d_o || !d_r || !d_t) return fail(GRLCU_ERR_NOMEM, "grlcu_model_step: out of device memory");
  CU(cudaMemcpyAsync(d_s, state, sizeof(double) * n * S, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemcpyAsync(d_a, action, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
  const int grd = grid_for(n, 128);
#define CALL(D) k_model_step<D><<<grd, 128, 0, c->stream>>>(n, c->p.tau, c->p.isteps, c->p.timeout, c->p.end_stop_penalty, d_s, d_a, d_n, d_o, d_r, d_t)
  DISPATCH_DYN(c, CALL);
#undef CALL
  c->launches++;
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(next, d_n, sizeof(double) * n * S, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(obs, d_o, sizeof(double) * n * OBS, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(reward, d_r, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(terminal, d_t, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

// ---- one-agent views (L2-signature shims) --------------------------------------------------------
template <int DYN>
__global__ void k_task_start(int64_t n, int test, double randomization, double timeout, int end_stop_penalty, const double *u,
                             double *state, double *obs, int32_t *terminal)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x)
  {
    double x[S], o[OBS];
    int term;
    task_start<DYN>(u[t], randomization, test, x);
    observe<DYN>(x, timeout, o, term, end_stop_penalty);
#pragma unroll
    for (int i = 0; i < S; ++i) state[t * S + i] = x[i];
#pragma unroll
    for (int i = 0; i < OBS; ++i) obs[t * OBS + i] = o[i];
    if (terminal) terminal[t] = term;
  }
}

extern "C" int grlcu_task_start(grlcu_ctx *c, int64_t n, int32_t test, const double *u, double *state, double *obs,
                                int32_t *terminal)
{
  if (!c || n < 0 || !u || !state || !obs) return fail(GRLCU_ERR_ARG, "grlcu_task_start: bad arguments");
  if (n == 0) return 0;
  CU(cudaSetDevice(c->device));
  const int S = c->S, OBS = c->OBS;
  Scratch s;
  double *d_u = s.get<double>(n), *d_s = s.get<double>((size_t)n * S), *d_o = s.get<double>((size_t)n * OBS);
  int32_t *d_t = s.get<int32_t>(n);
  if (!d_u || !d_s || !d_o || !d_t) return fail(GRLCU_ERR_NOMEM, "grlcu_task_start: out of device memory");
  CU(cudaMemcpyAsync(d_u, u, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
  const int grd = grid_for(n, 128);
#define CALL(D) k_task_start<D><<<grd, 128, 0, c->stream>>>(n, test, c->p.randomization, c->p.timeout, c->p.end_stop_penalty, d_u, d_s, d_o, d_t)
  DISPATCH_DYN(c, CALL);
#undef CALL
  c->launches++;
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(state, d_s, sizeof(double) * n * S, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(obs, d_o, sizeof(double) * n * OBS, cudaMemcpyDeviceToHost, c->stream));
  if (terminal) CU(cudaMemcpyAsync(terminal, d_t, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

// one thread: the reference's sequential order is the contract (linear.cpp:147-151, 207-215)
// [lo, hi]: the representation's output_min / output_max (+-DBL_MAX when unset, linear.cpp:88-98); the read is
// clamped to it (linear.cpp:179-180) and, with limit = 1, so is every updated weight (linear.cpp:207-211)
// perm: logical index -> physical slot (struct Geo), or NULL
__global__ void k_table_read(const double *tbl, const uint32_t *perm, int shift, int n_idx, const uint64_t *idx, uint64_t memory, double lo, double hi,
                             double *value, int *bad)
{
  double r = 0;
  for (int i = 0; i < n_idx; ++i)
  {
    if (idx[i] >= memory) { *bad = 1; return; }
    r += tbl[perm ? perm[idx[i]] : idx[i] << shift];
  }
  *value = fmin(fmax(r / n_idx, lo), hi);
}

__global__ void k_table_update(double *tbl, const uint32_t *perm, int shift, int n_idx, const uint64_t *idx, uint64_t memory, double lo, double hi,
                               double delta, int *bad)
{
  for (int i = 0; i < n_idx; ++i)
  {
    if (idx[i] == ~0ull) continue;                    // IndexProjection::invalid_index()
    if (idx[i] >= memory) { *bad = 1; return; }
    const uint64_t s = perm ? perm[idx[i]] : idx[i] << shift;
    tbl[s] = fmin(fmax(tbl[s] + delta, lo), hi);
  }
}

static int table_view(grlcu_ctx *c, bool actor, int32_t agent, int32_t n_idx, const uint64_t *idx, double *value, double delta)
{
  if (!c || agent < 0 || agent >= c->p.n || n_idx < 1 || !idx) return fail(GRLCU_ERR_ARG, "grlcu_read/update: bad arguments");
  if (actor && c->cfg.algo != GRLCU_ALGO_AC) return fail(GRLCU_ERR_UNSUPPORTED, "no actor table: not an actor-critic context");
  CU(cudaSetDevice(c->device));
  if (int r = ensure_plain(c)) return r;
  Scratch s;
  uint64_t *d_i = s.get<uint64_t>(n_idx);
  double *d_v = s.get<double>(1);
  int *d_bad = s.get<int>(1);
  if (!d_i || !d_v || !d_bad) return fail(GRLCU_ERR_NOMEM, "grlcu_read/update: out of device memory");
  CU(cudaMemcpyAsync(d_i, idx, sizeof(uint64_t) * n_idx, cudaMemcpyHostToDevice, c->stream));
  CU(cudaMemsetAsync(d_bad, 0, sizeof(int), c->stream));
  double *tbl = (actor ? c->ac.actor : c->p.table) + (size_t)agent * c->p.tstride;
  const int shift = c->ac.actor ? 1 : 0;          // actor-critic: pair x at doubles [2x, 2x + 1]
  const double lo = actor ? c->ac.action_min : -DBL_MAX, hi = actor ? c->ac.action_max : DBL_MAX;
  const uint32_t *perm = c->geo ? c->p.geo.perm : nullptr;
  if (value) k_table_read<<<1, 1, 0, c->stream>>>(tbl, perm, shift, n_idx, d_i, (uint64_t)c->cfg.memory, lo, hi, d_v, d_bad);
  else k_table_update<<<1, 1, 0, c->stream>>>(tbl, perm, shift, n_idx, d_i, (uint64_t)c->cfg.memory, lo, hi, delta, d_bad);
  c->launches++;
  CU(cudaGetLastError());
  int bad = 0;
  CU(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  if (value) CU(cudaMemcpyAsync(value, d_v, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (bad) return fail(GRLCU_ERR_ARG, "grlcu_read/update: index beyond memory (%d entries)", c->cfg.memory);
  return 0;
}

extern "C" int grlcu_read(grlcu_ctx *c, int32_t agent, int32_t n_idx, const uint64_t *idx, double *value)
{
  if (!value) return fail(GRLCU_ERR_ARG, "grlcu_read: null output");
  return table_view(c, false, agent, n_idx, idx, value, 0.);
}

extern "C" int grlcu_update(grlcu_ctx *c, int32_t agent, int32_t n_idx, const uint64_t *idx, double delta)
{
  return table_view(c, false, agent, n_idx, idx, nullptr, delta);
}

extern "C" int grlcu_read_actor(grlcu_ctx *c, int32_t agent, int32_t n_idx, const uint64_t *idx, double *value)
{
  if (!value) return fail(GRLCU_ERR_ARG, "grlcu_read_actor: null output");
  return table_view(c, true, agent, n_idx, idx, value, 0.);
}

extern "C" int grlcu_update_actor(grlcu_ctx *c, int32_t agent, int32_t n_idx, const uint64_t *idx, double delta)
{
  return table_view(c, true, agent, n_idx, idx, nullptr, delta);
}

// ---- fused run -------------------------------------------------------------------------------
static cudaEvent_t get_event(grlcu_ctx *c)
{
  if (!c->ev_pool.empty()) { cudaEvent_t e = c->ev_pool.back(); c->ev_pool.pop_back(); return e; }
  cudaEvent_t e;
  cudaEventCreate(&e);
  return e;
}

// Fold the event pairs of finished launches into the running total so that ev_pending stays short when a caller
// never asks for the kernel time (OnlineLearningExperiment::run loops grlcu_run + grlcu_get_results for hours).
static void drain_events(grlcu_ctx *c)
{
  size_t k = 0;
  for (auto &pr : c->ev_pending)
  {
    float t = 0;
    if (cudaEventQuery(pr.second) == cudaSuccess && cudaEventElapsedTime(&t, pr.first, pr.second) == cudaSuccess)
    {
      c->ev_ms += t; c->ev_n += 1;
      c->ev_pool.push_back(pr.first); c->ev_pool.push_back(pr.second);
    }
    else c->ev_pending[k++] = pr;
  }
  c->ev_pending.resize(k);
  cudaGetLastError();      // cudaEventQuery reports cudaErrorNotReady through the sticky-free error state
}

extern "C" int grlcu_run(grlcu_ctx *c, int64_t n_steps, int32_t cohort)
{
  if (!c || n_steps < 0) return fail(GRLCU_ERR_ARG, "grlcu_run: bad arguments");
  if (n_steps == 0) return 0;
  CU(cudaSetDevice(c->device));
  const int n = c->p.n;
  if (cohort <= 0 || cohort > n) cohort = n;
  // the cache stays live between launches (values return to the tables lazily, see ensure_plain) unless the
  // agents run in cohorts or GRLCU_EAGER_CLOSE is set
  if (c->ev_pending.size() >= 32) drain_events(c);
  c->traces_live = true;
  // the trace value cache applies to the replacing trace (unique owner slot per index); it stays live between
  // launches, agent by agent (values return to the tables lazily, see ensure_plain), unless GRLCU_EAGER_CLOSE is set
  const bool cached = c->p.mode == 0 && !c->p.force_slow && c->p.trace_kind == GRLCU_TRACE_REPLACING && c->p.trace_cap <= 64 &&
                      (c->cfg.algo == GRLCU_ALGO_AC || c->cfg.n_actions == 3);
  if (!cached && c->p.mode == 0 && c->p.trace_kind == GRLCU_TRACE_REPLACING && c->p.trace_cap > 64 && !c->warned_slow)
  {
    // (the fast update paths carry a 64-bit per-entry duplicate mask)
    fprintf(stderr, "libgrlcu: gamma*lambda = %g needs a %d-entry trace (> 64): this context runs the sequential update path, "
            "about 3x slower than the cached one\n", c->cfg.gamma * c->cfg.lambda, c->p.trace_cap);
    c->warned_slow = true;
  }
  const int do_close = getenv("GRLCU_EAGER_CLOSE") ? 1 : 0;
  // Register budget: a launch with more than 8 warps per SM gets the 128-register instantiation, which spills
  // ~0.5 KB per thread; two launches of half the agents at up to 255 registers are faster (profiles/r02_residency.txt:
  // 5.76e8 against 4.38e8 agent-steps/s for the same 32,768 agents).  So, unless the caller chose a cohort size, the
  // agents run in the fewest equal cohorts that fit 8 warps per SM.  Agents are independent: the order cannot matter.
  // (per-agent lambda: trace lengths differ from warp to warp; cohorts in agent-id order lose to one launch that backfills
  // blocks as the short-trace warps finish (3.27e8 against 3.97e8 on the sweep workload), cohorts in lambda order
  // (chunk_map, grlcu_set_hyper) win: 4.23e8 against 4.07e8, profiles/r02_sweep_lambda_sorted_cohorts.txt)
  if (cohort == n && (int)c->p.cblock < n && (!c->p.lambda_a || c->p.chunk_map) && !getenv("GRLCU_MINB7")) cohort = (int)c->p.cblock;
  for (int begin = 0; begin < n; begin += cohort)
  {
    const int end = begin + cohort < n ? begin + cohort : n;
    const int grd = (end - begin + c->block - 1) / c->block;
    cudaEvent_t e0 = get_event(c), e1 = get_event(c);
    CU(cudaEventRecord(e0, c->stream));
    // at most 8 warps per SM: the 255-register instantiation (see k_run)
    const int ctas_per_sm = (grd + c->sm_count - 1) / c->sm_count;
    const bool roomy = ctas_per_sm * (c->block / 32) <= 8 && !getenv("GRLCU_MINB7");
#define CALL_G(D, G)                                                                               \
    do {                                                                                           \
      if (cached && roomy) k_run<D, 16, 3, true, 4, G><<<grd, c->block, 0, c->stream>>>(c->p, begin, end, n_steps, do_close); \
      else if (cached) k_run<D, 16, 3, true, 7, G><<<grd, c->block, 0, c->stream>>>(c->p, begin, end, n_steps, do_close); \
      else if (roomy) k_run<D, 16, 3, false, 4, G><<<grd, c->block, 0, c->stream>>>(c->p, begin, end, n_steps, 0); \
      else k_run<D, 16, 3, false, 7, G><<<grd, c->block, 0, c->stream>>>(c->p, begin, end, n_steps, 0);      \
    } while (0)
#define CALL(D) CALL_G(D, false)
#define CALL_AC(D)                                                                                 \
    do {                                                                                           \
      if (cached && roomy) k_run_ac<D, 16, true, 4><<<grd, c->block, 0, c->stream>>>(c->p, c->ac, begin, end, n_steps, do_close); \
      else if (cached) k_run_ac<D, 16, true, 7><<<grd, c->block, 0, c->stream>>>(c->p, c->ac, begin, end, n_steps, do_close); \
      else if (roomy) k_run_ac<D, 16, false, 4><<<grd, c->block, 0, c->stream>>>(c->p, c->ac, begin, end, n_steps, 0); \
      else k_run_ac<D, 16, false, 7><<<grd, c->block, 0, c->stream>>>(c->p, c->ac, begin, end, n_steps, 0);      \
    } while (0)
#define CALL_SAFE(D) k_run<D, 16, 3, false, 7, false, true><<<grd, c->block, 0, c->stream>>>(c->p, begin, end, n_steps, 0)
    if (c->cfg.algo == GRLCU_ALGO_AC) DISPATCH_DYN(c, CALL_AC);
    else if (c->p.safe > 0) DISPATCH_DYN(c, CALL_SAFE);
    else if (c->cfg.n_actions != 3)
    {
#define CALL_A2(D) k_run<D, 16, 2, false, 7, false><<<grd, c->block, 0, c->stream>>>(c->p, begin, end, n_steps, 0)
#define CALL_A4(D) k_run<D, 16, 4, false, 7, false><<<grd, c->block, 0, c->stream>>>(c->p, begin, end, n_steps, 0)
#define CALL_A5(D) k_run<D, 16, 5, false, 7, false><<<grd, c->block, 0, c->stream>>>(c->p, begin, end, n_steps, 0)
      if (c->cfg.n_actions == 2) DISPATCH_DYN(c, CALL_A2);
      else if (c->cfg.n_actions == 4) DISPATCH_DYN(c, CALL_A4);
      else DISPATCH_DYN(c, CALL_A5);
#undef CALL_A2
#undef CALL_A4
#undef CALL_A5
    }
#define CALL_GEO(D) CALL_G(D, true)
    else if (c->geo) DISPATCH_DYN(c, CALL_GEO);    // tile-major (pendulum: from tile coordinates) or visit-ordered (perm only)
#undef CALL_GEO
    else DISPATCH_SARSA(c, CALL);
#undef CALL_SAFE
#undef CALL_AC
#undef CALL
#undef CALL_G
    if (cached && !do_close) c->cache_live = true;
    CU(cudaEventRecord(e1, c->stream));
    c->ev_pending.push_back({e0, e1});
    c->launches++;
  }
  CU(cudaGetLastError());
  return 0;
}

extern "C" int grlcu_sync(grlcu_ctx *c)
{
  if (!c) return fail(GRLCU_ERR_ARG, "null context");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int grlcu_run_kernel_time(grlcu_ctx *c, double *ms, int64_t *launches)
{
  if (!c) return fail(GRLCU_ERR_ARG, "null context");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  double total = c->ev_ms;
  for (auto &pr : c->ev_pending)
  {
    float t = 0;
    CU(cudaEventElapsedTime(&t, pr.first, pr.second));
    total += t;
    c->ev_pool.push_back(pr.first); c->ev_pool.push_back(pr.second);
  }
  if (ms) *ms = total;
  if (launches) *launches = c->ev_n + (int64_t)c->ev_pending.size();
  c->ev_pending.clear();
  c->ev_ms = 0; c->ev_n = 0;
  return 0;
}

extern "C" int grlcu_layout_info(grlcu_ctx *c, int32_t *tile_major, uint32_t *table_stride, uint32_t *inbox_words, uint32_t *shared_words)
{
  if (!c) return fail(GRLCU_ERR_ARG, "null context");
  if (tile_major) *tile_major = c->geo ? c->p.geo.enabled : 0;     // 1: from tile coordinates, 2: visit-ordered (perm only)
  if (table_stride) *table_stride = c->p.tstride;
  if (inbox_words) *inbox_words = c->inbox_words;
  if (shared_words) *shared_words = c->shared_words;
  return 0;
}

extern "C" int grlcu_get_pressure(grlcu_ctx *c, uint32_t *out)
{
  if (!c || !out) return fail(GRLCU_ERR_ARG, "grlcu_get_pressure: bad arguments");
  if (!c->p.pressure) return fail(GRLCU_ERR_ARG, "grlcu_get_pressure: the context has no claim table (safe = 0)");
  CU(cudaSetDevice(c->device));
  CU(cudaMemcpyAsync(out, c->p.pressure, sizeof(uint32_t) * c->p.n, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int64_t grlcu_launch_count(grlcu_ctx *c) { return c ? c->launches : -1; }
extern "C" int grlcu_trace_capacity(grlcu_ctx *c) { return c ? c->p.trace_cap : -1; }

// ---- results / tables --------------------------------------------------------------------------
template <typename Tp>
static int d2h(grlcu_ctx *c, Tp *dst, const Tp *src, size_t count)
{
  if (!dst) return 0;
  CU(cudaMemcpyAsync(dst, src, sizeof(Tp) * count, cudaMemcpyDeviceToHost, c->stream));
  return 0;
}

static int results_async(grlcu_ctx *c, grlcu_results *r)
{
  if (!r) return 0;
  const Params &p = c->p;
  const size_t n = (size_t)p.n;
  if (int e = d2h(c, r->last_return, p.last_return, n)) return e;
  if (int e = d2h(c, r->last_test_return, p.last_test_return, n)) return e;
  if (int e = d2h(c, r->return_sum, p.return_sum, n)) return e;
  if (int e = d2h(c, r->steps, p.ss, n)) return e;
  if (int e = d2h(c, r->episodes, p.episodes, n)) return e;
  if (int e = d2h(c, r->swept, p.swept, n)) return e;
  if (int e = d2h(c, r->table_writes, p.tblw, n)) return e;
  return 0;
}

extern "C" int grlcu_get_results(grlcu_ctx *c, grlcu_results *r)
{
  if (!c || !r) return fail(GRLCU_ERR_ARG, "grlcu_get_results: bad arguments");
  CU(cudaSetDevice(c->device));
  if (int e = results_async(c, r)) return e;
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

__global__ void k_transpose_start(const double *rows, double *soa, int n, int S)
{
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n) return;
  for (int i = 0; i < S; ++i) soa[(size_t)i * n + a] = rows[(size_t)a * S + i];
}

extern "C" int grlcu_run_host(grlcu_ctx *c, const double *start_states, int64_t n_steps, int32_t cohort,
                              grlcu_results *results)
{
  if (!c || n_steps < 0) return fail(GRLCU_ERR_ARG, "grlcu_run_host: bad arguments");
  CU(cudaSetDevice(c->device));
  const int n = c->p.n, S = c->S;
  if (start_states)
  {
    if (!c->d_start) if (int r = dalloc(c, &c->d_start, (size_t)n * S * 2)) return r;
    double *rows = c->d_start + (size_t)n * S;
    CU(cudaMemcpyAsync(rows, start_states, sizeof(double) * n * S, cudaMemcpyHostToDevice, c->stream));
    const int blk = 256, grd = (n + blk - 1) / blk;
    k_transpose_start<<<grd, blk, 0, c->stream>>>(rows, c->d_start, n, S);
    k_begin_override<<<grd, blk, 0, c->stream>>>(c->p);
    c->launches += 2;
    c->p.start_override = c->d_start;
  }
  if (int r = grlcu_run(c, n_steps, cohort)) return r;
  if (int r = results_async(c, results)) return r;
  CU(cudaStreamSynchronize(c->stream));
  CU(cudaGetLastError());
  return 0;
}

extern "C" int grlcu_get_table(grlcu_ctx *c, int32_t agent, double *out)
{
  if (!c || !out || agent < 0 || agent >= c->p.n) return fail(GRLCU_ERR_ARG, "grlcu_get_table: bad arguments");
  CU(cudaSetDevice(c->device));
  if (int r = ensure_plain(c)) return r;
  const double *src = c->p.table + (size_t)agent * c->p.tstride;
  if (c->ac.actor)
  {
    k_table_gather_pairs<<<grid_for(c->cfg.memory, 256), 256, 0, c->stream>>>(src, c->cfg.memory, c->d_stage);
    c->launches++;
    CU(cudaGetLastError());
    src = c->d_stage;
  }
  else if (c->geo)
  {
    // physical (tile-major) -> the reference's logical order
    k_table_gather<<<grid_for(c->cfg.memory, 256), 256, 0, c->stream>>>(src, c->p.geo.perm, c->cfg.memory, c->d_stage);
    c->launches++;
    CU(cudaGetLastError());
    src = c->d_stage;
  }
  CU(cudaMemcpyAsync(out, src, sizeof(double) * c->cfg.memory, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int grlcu_set_table(grlcu_ctx *c, int32_t agent, const double *in)
{
  if (!c || !in || agent < 0 || agent >= c->p.n) return fail(GRLCU_ERR_ARG, "grlcu_set_table: bad arguments");
  CU(cudaSetDevice(c->device));
  if (int r = ensure_plain(c)) return r;
  // The fused kernel keeps a directory inside the table: a word whose high half is kSentinelHi (a signalling-NaN
  // pattern no arithmetic produces) names a cache slot.  A checkpoint must not smuggle one in.
  for (int32_t x = 0; x < c->cfg.memory; ++x)
  {
    uint64_t bits;
    memcpy(&bits, in + x, sizeof(bits));
    if ((uint32_t)(bits >> 32) == (uint32_t)kSentinelHi)
      return fail(GRLCU_ERR_ARG, "grlcu_set_table: entry %d holds the NaN pattern 0x%08X........ that the kernels reserve", x, kSentinelHi);
  }
  double *dst = c->p.table + (size_t)agent * c->p.tstride;
  if (c->ac.actor)
  {
    CU(cudaMemcpyAsync(c->d_stage, in, sizeof(double) * c->cfg.memory, cudaMemcpyHostToDevice, c->stream));
    k_table_scatter_pairs<<<grid_for(c->cfg.memory, 256), 256, 0, c->stream>>>(dst, c->cfg.memory, c->d_stage);
    c->launches++;
    CU(cudaGetLastError());
  }
  else if (c->geo)
  {
    CU(cudaMemcpyAsync(c->d_stage, in, sizeof(double) * c->cfg.memory, cudaMemcpyHostToDevice, c->stream));
    k_table_scatter<<<grid_for(c->cfg.memory, 256), 256, 0, c->stream>>>(dst, c->p.geo.perm, c->cfg.memory, c->d_stage);
    c->launches++;
    CU(cudaGetLastError());
  }
  else
    CU(cudaMemcpyAsync(dst, in, sizeof(double) * c->cfg.memory, cudaMemcpyHostToDevice, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int grlcu_get_actor_table(grlcu_ctx *c, int32_t agent, double *out)
{
  if (!c || !out || agent < 0 || agent >= c->p.n) return fail(GRLCU_ERR_ARG, "grlcu_get_actor_table: bad arguments");
  if (!c->ac.actor) return fail(GRLCU_ERR_ARG, "grlcu_get_actor_table: not an actor-critic context");
  CU(cudaSetDevice(c->device));
  k_table_gather_pairs<<<grid_for(c->cfg.memory, 256), 256, 0, c->stream>>>(c->ac.actor + (size_t)agent * c->p.tstride, c->cfg.memory, c->d_stage);
  c->launches++;
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(out, c->d_stage, sizeof(double) * c->cfg.memory, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int grlcu_set_actor_table(grlcu_ctx *c, int32_t agent, const double *in)
{
  if (!c || !in || agent < 0 || agent >= c->p.n) return fail(GRLCU_ERR_ARG, "grlcu_set_actor_table: bad arguments");
  if (!c->ac.actor) return fail(GRLCU_ERR_ARG, "grlcu_set_actor_table: not an actor-critic context");
  CU(cudaSetDevice(c->device));
  CU(cudaMemcpyAsync(c->d_stage, in, sizeof(double) * c->cfg.memory, cudaMemcpyHostToDevice, c->stream));
  k_table_scatter_pairs<<<grid_for(c->cfg.memory, 256), 256, 0, c->stream>>>(c->ac.actor + (size_t)agent * c->p.tstride, c->cfg.memory, c->d_stage);
  c->launches++;
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int grlcu_get_state(grlcu_ctx *c, int32_t agent, double *state, double *action, int32_t *in_episode)
{
  if (!c || agent < 0 || agent >= c->p.n) return fail(GRLCU_ERR_ARG, "grlcu_get_state: bad arguments");
  CU(cudaSetDevice(c->device));
  CU(cudaStreamSynchronize(c->stream));
  if (state)
    for (int i = 0; i < c->S; ++i)
      CU(cudaMemcpy(state + i, c->p.state + (size_t)i * c->p.n + agent, sizeof(double), cudaMemcpyDeviceToHost));
  if (action) CU(cudaMemcpy(action, c->p.action + agent, sizeof(double), cudaMemcpyDeviceToHost));
  if (in_episode)
  {
    int32_t f = 0;
    CU(cudaMemcpy(&f, c->p.flags + agent, sizeof(int32_t), cudaMemcpyDeviceToHost));
    *in_episode = f & 1;
  }
  return 0;
}

// ---- replay ------------------------------------------------------------------------------------
extern "C" int grlcu_replay_step(grlcu_ctx *c, const grlcu_replay_in *in, grlcu_replay_out *out)
{
  if (!c || !in || !out) return fail(GRLCU_ERR_ARG, "grlcu_replay_step: bad arguments");
  const bool is_ac = c->cfg.algo == GRLCU_ALGO_AC;
  if (!in->started || !in->state || !in->action || !in->test || !in->obs || !in->reward || !in->terminal ||
      !in->u || !in->start_state || !in->start_u ||
      (!is_ac && (!in->r_explore || !in->r_tie || !in->start_r_explore || !in->start_r_tie)))
    return fail(GRLCU_ERR_ARG, "grlcu_replay_step: missing input array");
  CU(cudaSetDevice(c->device));
  if (int r = ensure_plain(c)) return r;
  c->traces_live = true;
  // AC mode: one projection row (the fed observation) and four value rows (pi(s'), V(s'), u, a')
  const size_t n = (size_t)c->p.n, S = (size_t)c->S, OBS = (size_t)c->OBS, T = (size_t)c->cfg.tilings,
               A = is_ac ? 4 : (size_t)c->cfg.n_actions, HA = is_ac ? 1 : A, cap = (size_t)c->p.trace_cap;
  Scratch s;
  ReplayIO io;
  memset(&io, 0, sizeof(io));
  bool oom = false;
#define UP(field, type, count)                                                                  \
  do {                                                                                          \
    type *d = s.get<type>(count);                                                               \
    if (!d) { oom = true; break; }                                                              \
    if (cudaMemcpyAsync(d, in->field, sizeof(type) * (count), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) oom = true; \
    io.field = d;                                                                               \
  } while (0)
  UP(started, int32_t, n); UP(start_state, double, S * n); UP(start_u, double, n);
  if (!is_ac) { UP(start_r_explore, uint32_t, n); UP(start_r_tie, uint32_t, n); }
  UP(state, double, S * n); UP(action, double, n); UP(test, int32_t, n);
  UP(obs, double, OBS * n); UP(reward, double, n); UP(terminal, int32_t, n);
  UP(u, double, n);
  if (!is_ac) { UP(r_explore, uint32_t, n); UP(r_tie, uint32_t, n); }
  if (in->force_action && !is_ac) UP(force_action, int32_t, n);
#undef UP
#define OUTB(ofield, field, type, count) \
  do { if (out->field) { io.ofield = s.get<type>(count); if (!io.ofield) oom = true; } } while (0)
  OUTB(o_next, next, double, S * n); OUTB(o_obs, obs, double, OBS * n); OUTB(o_reward, reward, double, n);
  OUTB(o_terminal, terminal, int32_t, n); OUTB(o_hash, hash, uint32_t, HA * T * n);
  OUTB(o_index, index, uint32_t, HA * T * n); OUTB(o_q, q, double, A * n);
  OUTB(o_action_index, action_index, int32_t, n); OUTB(o_start_q, start_q, double, A * n);
  OUTB(o_start_action_index, start_action_index, int32_t, n);
  OUTB(o_q_prev, q_prev, double, n); OUTB(o_target, target, double, n); OUTB(o_delta, delta, double, n);
  OUTB(o_swept, swept, int32_t, n); OUTB(o_trace_len, trace_len, int32_t, n);
  OUTB(o_trace_index, trace_index, uint32_t, cap * T * n);
#undef OUTB
  if (oom) return fail(GRLCU_ERR_NOMEM, "grlcu_replay_step: device scratch allocation/copy failed");
  const int blk = 64, grd = ((int)n + blk - 1) / blk;
#define CALL(D) k_replay<D, 16, 3, false><<<grd, blk, 0, c->stream>>>(c->p, io)
#define CALL_AC(D) k_replay_ac<D, 16><<<grd, blk, 0, c->stream>>>(c->p, c->ac, io)
#define CALL_SAFE(D) k_replay<D, 16, 3, false, true><<<grd, blk, 0, c->stream>>>(c->p, io)
  if (is_ac) DISPATCH_DYN(c, CALL_AC);
  else if (c->p.safe > 0) DISPATCH_DYN(c, CALL_SAFE);
  else if (c->cfg.n_actions != 3)
  {
#define CALL_A2(D) k_replay<D, 16, 2, false><<<grd, blk, 0, c->stream>>>(c->p, io)
#define CALL_A4(D) k_replay<D, 16, 4, false><<<grd, blk, 0, c->stream>>>(c->p, io)
#define CALL_A5(D) k_replay<D, 16, 5, false><<<grd, blk, 0, c->stream>>>(c->p, io)
    if (c->cfg.n_actions == 2) DISPATCH_DYN(c, CALL_A2);
    else if (c->cfg.n_actions == 4) DISPATCH_DYN(c, CALL_A4);
    else DISPATCH_DYN(c, CALL_A5);
#undef CALL_A2
#undef CALL_A4
#undef CALL_A5
  }
#define CALL_GEO(D) k_replay<D, 16, 3, true><<<grd, blk, 0, c->stream>>>(c->p, io)
  else if (c->geo) DISPATCH_DYN(c, CALL_GEO);
#undef CALL_GEO
  else DISPATCH_SARSA(c, CALL);
#undef CALL_AC
#undef CALL
  c->launches++;
  CU(cudaGetLastError());
#define DOWN(ofield, field, type, count) \
  do { if (out->field) CU(cudaMemcpyAsync(out->field, io.ofield, sizeof(type) * (count), cudaMemcpyDeviceToHost, c->stream)); } while (0)
  DOWN(o_next, next, double, S * n); DOWN(o_obs, obs, double, OBS * n); DOWN(o_reward, reward, double, n);
  DOWN(o_terminal, terminal, int32_t, n); DOWN(o_hash, hash, uint32_t, HA * T * n);
  DOWN(o_index, index, uint32_t, HA * T * n); DOWN(o_q, q, double, A * n);
  DOWN(o_action_index, action_index, int32_t, n); DOWN(o_start_q, start_q, double, A * n);
  DOWN(o_start_action_index, start_action_index, int32_t, n);
  DOWN(o_q_prev, q_prev, double, n); DOWN(o_target, target, double, n); DOWN(o_delta, delta, double, n);
  DOWN(o_swept, swept, int32_t, n); DOWN(o_trace_len, trace_len, int32_t, n);
  DOWN(o_trace_index, trace_index, uint32_t, cap * T * n);
#undef DOWN
  int32_t mismatch = 0, oob = 0;
  if (c->geo) CU(cudaMemcpyAsync(&mismatch, c->p.geo.mismatch, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(&oob, c->p.oob, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  if (oob) return fail(GRLCU_ERR_CUDA, "grlcu_replay_step: a physical table slot beyond the agent's row (%u doubles) was computed", c->p.tstride);
  if (mismatch)
    return fail(GRLCU_ERR_CUDA, "grlcu_replay_step: a tile's geometric slot differs from perm[hash index] (layout bug)");
  return 0;
}

// ---- several GPUs of one box (include/grlcu.h) ------------------------------------------------------------------
namespace {
struct NcclApi
{
  void *handle = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};

// By soname: a process that already holds an NCCL (torch's bundled one) gets that one, not a second copy.
int load_nccl(NcclApi *api)
{
  static NcclApi g;
  if (!g.handle)
  {
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) return fail(GRLCU_ERR_UNSUPPORTED, "grlcu_group: cannot load libnccl.so.2 (%s)", dlerror());
    g.CommInitAll = (decltype(g.CommInitAll))dlsym(h, "ncclCommInitAll");
    g.CommDestroy = (decltype(g.CommDestroy))dlsym(h, "ncclCommDestroy");
    g.GroupStart = (decltype(g.GroupStart))dlsym(h, "ncclGroupStart");
    g.GroupEnd = (decltype(g.GroupEnd))dlsym(h, "ncclGroupEnd");
    g.AllGather = (decltype(g.AllGather))dlsym(h, "ncclAllGather");
    g.GetErrorString = (decltype(g.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!g.CommInitAll || !g.CommDestroy || !g.GroupStart || !g.GroupEnd || !g.AllGather || !g.GetErrorString)
    {
      dlclose(h);
      return fail(GRLCU_ERR_UNSUPPORTED, "grlcu_group: libnccl.so.2 lacks a needed entry point");
    }
    g.handle = h;
  }
  *api = g;
  return 0;
}
}  // namespace

extern "C" int grlcu_device_count(void)
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

constexpr int kResultWords = 7;    // 8-byte words per agent in the gather: the seven grlcu_results fields

struct grlcu_group
{
  std::vector<grlcu_ctx *> ctxs;
  std::vector<ncclComm_t> comms;
  std::vector<uint64_t *> send, recv;      // per context: [kResultWords][n] and [n_ctx][kResultWords][n]
  std::vector<uint64_t> host;              // pinned-size staging of one recv buffer
  NcclApi nccl;
  int n;                                   // agents per context
};

// the seven per-agent result arrays of one context as rows of 8-byte words (episodes widened)
__global__ void k_pack_results(Params p, uint64_t *out)
{
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= p.n) return;
  const size_t n = (size_t)p.n;
  out[0 * n + a] = (uint64_t)__double_as_longlong(p.last_return[a]);
  out[1 * n + a] = (uint64_t)__double_as_longlong(p.last_test_return[a]);
  out[2 * n + a] = (uint64_t)__double_as_longlong(p.return_sum[a]);
  out[3 * n + a] = p.ss[a];
  out[4 * n + a] = (uint64_t)p.episodes[a];
  out[5 * n + a] = p.swept[a];
  out[6 * n + a] = p.tblw[a];
}

extern "C" int grlcu_group_create(grlcu_ctx *const *ctxs, int32_t n_ctx, grlcu_group **out)
{
  if (!ctxs || n_ctx < 1 || !out) return fail(GRLCU_ERR_ARG, "grlcu_group_create: bad arguments");
  *out = nullptr;
  for (int i = 0; i < n_ctx; ++i)
  {
    if (!ctxs[i]) return fail(GRLCU_ERR_ARG, "grlcu_group_create: null context");
    if (ctxs[i]->p.n != ctxs[0]->p.n) return fail(GRLCU_ERR_ARG, "grlcu_group_create: every context must hold the same number of agents");
    for (int j = 0; j < i; ++j)
      if (ctxs[j]->device == ctxs[i]->device) return fail(GRLCU_ERR_ARG, "grlcu_group_create: two contexts on device %d", ctxs[i]->device);
  }
  grlcu_group *g = new (std::nothrow) grlcu_group();
  if (!g) return fail(GRLCU_ERR_NOMEM, "out of host memory");
  g->ctxs.assign(ctxs, ctxs + n_ctx);
  g->n = ctxs[0]->p.n;
  g->send.assign(n_ctx, nullptr); g->recv.assign(n_ctx, nullptr);
  const size_t words = (size_t)kResultWords * g->n;
  int r = 0;
  for (int i = 0; i < n_ctx && !r; ++i)
  {
    if (cudaSetDevice(ctxs[i]->device) != cudaSuccess) { r = fail(GRLCU_ERR_CUDA, "cudaSetDevice failed"); break; }
    if (cudaMalloc(&g->send[i], words * sizeof(uint64_t)) != cudaSuccess ||
        cudaMalloc(&g->recv[i], words * n_ctx * sizeof(uint64_t)) != cudaSuccess)
      r = fail(GRLCU_ERR_NOMEM, "grlcu_group_create: out of device memory");
  }
  if (!r && n_ctx > 1)
  {
    r = load_nccl(&g->nccl);
    if (!r)
    {
      std::vector<int> devs(n_ctx);
      for (int i = 0; i < n_ctx; ++i) devs[i] = ctxs[i]->device;
      g->comms.assign(n_ctx, nullptr);
      ncclResult_t e = g->nccl.CommInitAll(g->comms.data(), n_ctx, devs.data());
      if (e != ncclSuccess) { g->comms.clear(); r = fail(GRLCU_ERR_CUDA, "ncclCommInitAll failed: %s", g->nccl.GetErrorString(e)); }
    }
  }
  g->host.resize(words * n_ctx);
  if (r) { grlcu_group_destroy(g); return r; }
  *out = g;
  return 0;
}

extern "C" int grlcu_group_destroy(grlcu_group *g)
{
  if (!g) return 0;
  for (size_t i = 0; i < g->ctxs.size(); ++i)
  {
    cudaSetDevice(g->ctxs[i]->device);
    cudaStreamSynchronize(g->ctxs[i]->stream);
    if (i < g->comms.size() && g->comms[i]) g->nccl.CommDestroy(g->comms[i]);
    if (g->send[i]) cudaFree(g->send[i]);
    if (g->recv[i]) cudaFree(g->recv[i]);
  }
  delete g;
  return 0;
}

extern "C" int grlcu_group_run(grlcu_group *g, int64_t n_steps)
{
  if (!g) return fail(GRLCU_ERR_ARG, "null group");
  for (grlcu_ctx *c : g->ctxs)
    if (int r = grlcu_run(c, n_steps, 0)) return r;       // asynchronous: the devices run concurrently
  return 0;
}

extern "C" int grlcu_group_results(grlcu_group *g, grlcu_results *res, double *collective_ms)
{
  if (!g || !res) return fail(GRLCU_ERR_ARG, "grlcu_group_results: bad arguments");
  const int G = (int)g->ctxs.size(), n = g->n;
  const size_t words = (size_t)kResultWords * n;
  for (int i = 0; i < G; ++i)
  {
    grlcu_ctx *c = g->ctxs[i];
    CU(cudaSetDevice(c->device));
    k_pack_results<<<(n + 255) / 256, 256, 0, c->stream>>>(c->p, g->send[i]);
    c->launches++;
    CU(cudaGetLastError());
  }
  const uint64_t *src = g->send[0];
  if (collective_ms) *collective_ms = 0;
  if (G > 1)
  {
    // the job's only collective: every GPU receives every GPU's block (NVLink / NVSwitch)
    cudaEvent_t e0, e1;
    CU(cudaSetDevice(g->ctxs[0]->device));
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    for (int i = 0; i < G; ++i) { CU(cudaSetDevice(g->ctxs[i]->device)); CU(cudaStreamSynchronize(g->ctxs[i]->stream)); }
    CU(cudaSetDevice(g->ctxs[0]->device));
    CU(cudaEventRecord(e0, g->ctxs[0]->stream));
    ncclResult_t e = g->nccl.GroupStart();
    for (int i = 0; i < G && e == ncclSuccess; ++i)
    {
      CU(cudaSetDevice(g->ctxs[i]->device));
      e = g->nccl.AllGather(g->send[i], g->recv[i], words, ncclUint64, g->comms[i], g->ctxs[i]->stream);
    }
    if (e == ncclSuccess) e = g->nccl.GroupEnd();
    if (e != ncclSuccess) return fail(GRLCU_ERR_CUDA, "ncclAllGather failed: %s", g->nccl.GetErrorString(e));
    CU(cudaSetDevice(g->ctxs[0]->device));
    CU(cudaEventRecord(e1, g->ctxs[0]->stream));
    for (int i = 0; i < G; ++i) { CU(cudaSetDevice(g->ctxs[i]->device)); CU(cudaStreamSynchronize(g->ctxs[i]->stream)); }
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    if (collective_ms) *collective_ms = ms;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    src = g->recv[0];
  }
  CU(cudaSetDevice(g->ctxs[0]->device));
  CU(cudaMemcpyAsync(g->host.data(), src, words * G * sizeof(uint64_t), cudaMemcpyDeviceToHost, g->ctxs[0]->stream));
  CU(cudaStreamSynchronize(g->ctxs[0]->stream));
  for (int i = 0; i < G; ++i)
  {
    const uint64_t *blk = g->host.data() + (size_t)i * words;
    const size_t o = (size_t)i * n;
    auto f64 = [&](double *dst, int row) { if (dst) memcpy(dst + o, blk + (size_t)row * n, sizeof(double) * n); };
    auto u64 = [&](uint64_t *dst, int row) { if (dst) memcpy(dst + o, blk + (size_t)row * n, sizeof(uint64_t) * n); };
    f64(res->last_return, 0); f64(res->last_test_return, 1); f64(res->return_sum, 2);
    u64(res->steps, 3); u64(res->swept, 5); u64(res->table_writes, 6);
    if (res->episodes) for (int a = 0; a < n; ++a) res->episodes[o + a] = (uint32_t)blk[(size_t)4 * n + a];
  }
  return 0;
}

// ---- grlcu_project_noalloc (include/grlcu.h): the K2 projection without any allocation --------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  grlcu_project stages its inputs and
// outputs in device scratch that it cudaMallocs and frees per call; this variant stages them in ONE file-scope device
// array that every GPU loading the library holds until exit (768 KiB).  Layout of the array (offsets 256-byte aligned):
//   [0, 256 KiB)        inputs   double   [GRLCU_NOALLOC_MAX_INPUTS][GRLCU_MAX_DIMS]  (n * dims used)
//   [256, 512 KiB)      hashes   uint32_t [GRLCU_NOALLOC_MAX_INPUTS][16]
//   [512, 768 KiB)      indices  uint32_t [GRLCU_NOALLOC_MAX_INPUTS][16]
// Every call shares the array and it keeps the last call's contents.  Nothing in it has to start at zero: a call
// overwrites the n * dims inputs it reads and the kernel writes every one of the n * T hashes and indices that are copied
// out; words beyond those are never read.  Calls that overlap on one GPU (two streams, two host threads, two contexts on
// the same device) RACE on the array: the caller serialises them.
static __device__ __align__(256) unsigned char g_noalloc_scratch[GRLCU_NOALLOC_MAX_INPUTS * (GRLCU_MAX_DIMS * 8 + 16 * 4 + 16 * 4)];

namespace {
constexpr int kNoAllocMaxDevices = 64;
// In host code `g_noalloc_scratch` is a host-side stand-in, NOT a device address: the address the kernel gets comes from
// cudaGetSymbolAddress, looked up once per device with that device current.
unsigned char *g_noalloc_ptr[kNoAllocMaxDevices];
bool g_noalloc_have[kNoAllocMaxDevices];

int noalloc_scratch(int device, unsigned char **out)
{
  if (device < 0 || device >= kNoAllocMaxDevices) return fail(GRLCU_ERR_ARG, "grlcu_project_noalloc: device %d out of range", device);
  if (!g_noalloc_have[device])
  {
    CU(cudaSetDevice(device));
    void *ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr)
      return fail(GRLCU_ERR_CUDA, "grlcu_project_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
    g_noalloc_ptr[device] = (unsigned char *)ptr;
    g_noalloc_have[device] = true;
  }
  *out = g_noalloc_ptr[device];
  return 0;
}
}  // namespace

extern "C" int grlcu_project_noalloc(grlcu_ctx *c, int64_t n, const double *in, uint32_t *hash, uint32_t *index)
{
  if (!c || n < 0 || !in || !index) return fail(GRLCU_ERR_ARG, "grlcu_project_noalloc: bad arguments");
  if (n > GRLCU_NOALLOC_MAX_INPUTS)
    return fail(GRLCU_ERR_ARG, "grlcu_project_noalloc: %lld inputs exceed the static scratch of %d (GRLCU_NOALLOC_MAX_INPUTS): "
                "split the batch or use grlcu_project", (long long)n, GRLCU_NOALLOC_MAX_INPUTS);
  if (c->cfg.tilings != 16) return fail(GRLCU_ERR_UNSUPPORTED, "grlcu_project_noalloc: tilings %d (16 expected)", c->cfg.tilings);
  if (n == 0) return 0;
  CU(cudaSetDevice(c->device));
  unsigned char *base = nullptr;
  if (int r = noalloc_scratch(c->device, &base)) return r;
  const bool is_ac = c->cfg.algo == GRLCU_ALGO_AC;
  const int dims = is_ac ? c->cfg.obs_dims : c->cfg.obs_dims + 1, T = 16;
  double *d_in = (double *)base;
  uint32_t *d_h = (uint32_t *)(base + (size_t)GRLCU_NOALLOC_MAX_INPUTS * GRLCU_MAX_DIMS * 8);
  uint32_t *d_i = d_h + (size_t)GRLCU_NOALLOC_MAX_INPUTS * 16;
  CU(cudaMemcpyAsync(d_in, in, sizeof(double) * n * dims, cudaMemcpyHostToDevice, c->stream));
  const int grd = grid_for(n, 128);
#define CALL(D) k_project<Dyn<D>::OBS + 1, 16><<<grd, 128, 0, c->stream>>>(c->p.proj, n, d_in, d_h, d_i)
#define CALL_AC(D) k_project<Dyn<D>::OBS, 16><<<grd, 128, 0, c->stream>>>(c->p.proj, n, d_in, d_h, d_i)
  if (is_ac) DISPATCH_DYN(c, CALL_AC);
  else DISPATCH_SARSA(c, CALL);
#undef CALL_AC
#undef CALL
  c->launches++;
  CU(cudaGetLastError());
  if (hash) CU(cudaMemcpyAsync(hash, d_h, sizeof(uint32_t) * n * T, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaMemcpyAsync(index, d_i, sizeof(uint32_t) * n * T, cudaMemcpyDeviceToHost, c->stream));
  CU(cudaStreamSynchronize(c->stream));
  return 0;
}
