// grlcu_kernels.cuh -- __global__ kernels of the batched GRL learning loop (sm_100a).
#pragma once
#include "grlcu_agent.cuh"
#include "grlcu_cached.cuh"

namespace grlcu {

// K5: LinearRepresentation::reset (linear.cpp:104-125): params = init_min + U*(init_max-init_min),
// U from Philox(seed, agent) stream STREAM_TABLE, counter = entry pair.  One thread per entry pair,
// 16-byte stores, fully coalesced.
__global__ void k_table_init(double *table, int32_t n, int32_t memory, int64_t first_id, uint64_t seed,
                             double init_min, double init_max, int stream_id)
{
  const int64_t pairs = ((int64_t)memory + 1) / 2;
  const int64_t total = pairs * n;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x)
  {
    const int64_t a = t / pairs, pi = t - a * pairs;
    Philox4 w = philox_for(seed, (uint64_t)(first_id + a), stream_id, (uint64_t)pi);
    double v0 = init_min + u53(w.w[0], w.w[1]) * (init_max - init_min);
    double v1 = init_min + u53(w.w[2], w.w[3]) * (init_max - init_min);
    double *dst = table + a * (int64_t)memory + 2 * pi;
    if (2 * pi + 1 < memory && ((memory & 1) == 0))
      *reinterpret_cast<double2 *>(dst) = make_double2(v0, v1);
    else
    {
      dst[0] = v0;
      if (2 * pi + 1 < memory) dst[1] = v1;
    }
  }
}

// The same initial values under the tile-major physical layout (struct Geo): one thread per physical slot, the
// value of logical entry x = iperm[slot] is half (x & 1) of Philox counter x / 2; padding slots are zero.
__global__ void k_table_init_phys(double *table, int32_t n, uint32_t tstride, const uint32_t *iperm, int64_t first_id,
                                  uint64_t seed, double init_min, double init_max, int stream_id)
{
  const int64_t total = (int64_t)tstride * n;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x)
  {
    const int64_t a = t / tstride;
    const uint32_t x = __ldg(iperm + (t - a * tstride));
    double v = 0;
    if (x != kInvalidIndex)
    {
      Philox4 w = philox_for(seed, (uint64_t)(first_id + a), stream_id, (uint64_t)(x >> 1));
      v = init_min + ((x & 1u) ? u53(w.w[2], w.w[3]) : u53(w.w[0], w.w[1])) * (init_max - init_min);
    }
    table[t] = v;
  }
}

// Actor-critic: the interleaved {critic, actor} pairs of every agent (grlcu_ac.cuh): pair x = doubles [2x, 2x + 1] of the
// agent's row; critic values from Philox stream STREAM_TABLE, actor values from STREAM_ACTOR_TABLE, each exactly what
// k_table_init gives entry x (half x & 1 of counter x / 2).  One thread per pair, 16-byte coalesced stores.
__global__ void k_table_init_pairs(double *table, int32_t n, int32_t memory, int64_t first_id, uint64_t seed,
                                   double init_min, double init_max)
{
  const int64_t total = (int64_t)memory * n;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x)
  {
    const int64_t a = t / memory;
    const uint32_t x = (uint32_t)(t - a * memory);
    const Philox4 wc = philox_for(seed, (uint64_t)(first_id + a), STREAM_TABLE, (uint64_t)(x >> 1));
    const Philox4 wa = philox_for(seed, (uint64_t)(first_id + a), STREAM_ACTOR_TABLE, (uint64_t)(x >> 1));
    const double vc = init_min + ((x & 1u) ? u53(wc.w[2], wc.w[3]) : u53(wc.w[0], wc.w[1])) * (init_max - init_min);
    const double va = init_min + ((x & 1u) ? u53(wa.w[2], wa.w[3]) : u53(wa.w[0], wa.w[1])) * (init_max - init_min);
    *reinterpret_cast<double2 *>(table + 2 * t) = make_double2(vc, va);
  }
}

// strided copies of one table out of / into an agent's interleaved pairs (tbl = row base, + 1 for the actor)
__global__ void k_table_gather_pairs(const double *tbl, int32_t memory, double *out)
{
  for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < memory; x += gridDim.x * blockDim.x) out[x] = tbl[2 * (size_t)x];
}
__global__ void k_table_scatter_pairs(double *tbl, int32_t memory, const double *in)
{
  for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < memory; x += gridDim.x * blockDim.x) tbl[2 * (size_t)x] = in[x];
}

// logical <-> physical copies of one agent's table (grlcu_get_table / grlcu_set_table under the tile-major layout)
__global__ void k_table_gather(const double *tbl, const uint32_t *perm, int32_t memory, double *out)
{
  for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < memory; x += gridDim.x * blockDim.x) out[x] = tbl[__ldg(perm + x)];
}
__global__ void k_table_scatter(double *tbl, const uint32_t *perm, int32_t memory, const double *in)
{
  for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < memory; x += gridDim.x * blockDim.x) tbl[__ldg(perm + x)] = in[x];
}

// K2: TileCodingProjector::_project for n independent inputs ([n][DIMS] row-major).
template <int DIMS, int T>
__global__ void k_project(Projector proj, int64_t n, const double *in, uint32_t *hash, uint32_t *index)
{
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x)
  {
    double v[DIMS];
    int32_t q[DIMS];
#pragma unroll
    for (int i = 0; i < DIMS; ++i) v[i] = in[t * DIMS + i];
    quantize<DIMS, T>(proj, v, q);
#pragma unroll
    for (int j = 0; j < T; ++j)
    {
      uint32_t h = tile_hash_of<DIMS, T>(proj, q, j);
      if (hash) hash[t * T + j] = h;
      index[t * T + j] = h % (uint32_t)proj.memory;
    }
  }
}

// K1: ModeledEnvironment::step body for n independent (state, action) pairs (row-major).
template <int DYN>
__global__ void k_model_step(int64_t n, double tau, int isteps, double timeout, int end_stop_penalty,
                             const double *state, const double *action, double *next, double *obs, double *reward, int32_t *terminal)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x)
  {
    double x[S], x0[S], o[OBS];
    int term;
#pragma unroll
    for (int i = 0; i < S; ++i) x0[i] = x[i] = state[t * S + i];
    const double act = action[t];
    rk4<DYN>(tau, isteps, Dyn<DYN>::actuate(act), x);
    observe<DYN>(x, timeout, o, term, end_stop_penalty);
    double r = evaluate<DYN>(x0, act, x, end_stop_penalty);
#pragma unroll
    for (int i = 0; i < S; ++i) next[t * S + i] = x[i];
#pragma unroll
    for (int i = 0; i < OBS; ++i) obs[t * OBS + i] = o[i];
    reward[t] = r;
    terminal[t] = term;
  }
}

__device__ __forceinline__ Draws draws_from(const Philox4 &w)
{
  Draws d;
  d.u = u53(w.w[0], w.w[1]);
  d.r_explore = w.w[2] >> 1;
  d.r_tie = w.w[3] >> 1;
  return d;
}

// Draws of one act(): learning episodes use the agent's STREAM_ACT counter (registers), test
// episodes the STREAM_TEST_ACT counter kept in global memory (rare).
template <int DYN, int T>
__device__ __forceinline__ Draws next_draws(const Params &p, Agent<DYN, T> &ag, int a, uint64_t gid, bool test)
{
  if (test)
  {
    const uint64_t c = p.test_act_ctr[a];
    p.test_act_ctr[a] = c + 1;
    return draws_from(philox_for(p.seed, gid, STREAM_TEST_ACT, c));
  }
  return draws_from(philox_for(p.seed, gid, STREAM_ACT, ag.act_ctr++));
}

struct NoResolve { __device__ __forceinline__ double operator()(double v) const { return v; } };
template <int T> struct CacheResolve
{
  Cache<T> c;
  __device__ __forceinline__ double operator()(double v) const { return resolve<T>(c, v); }
};

// The agent's part of one loop iteration, written ONCE (the fused kernels are instruction-fetch bound, so the
// episode-start and the step/end paths share one copy of the hashing, the gathers and the update):
//   act   : there is a next action to choose (TDAgent::start / step; false for TDAgent::end, td.cpp:76-81)
//   learn : there is a transition to learn from (TDAgent::step / end of a learning episode, td.cpp:63-81)
// Returns the policy's own action choice (-1 when !act); commits `force` instead when force >= 0 (replay).
// SAFE (projector safe > 0, uncached kernels only): the policy looks its locations up in the agent's claim table at act
// time (claiming only with safe > 1); the predictor then projects (prev_obs, prev_action) and (obs, action) AGAIN, claiming
// both (sarsa.cpp:102-108 through Projector::project(base, variant), tile_coding.h:62-65), and reads Q(s', a') and Q(p) at
// those locations -- which can differ from the act-time ones once a claim lands on a probe path.
template <int DYN, int T, int A, bool CACHED, bool GEO, bool SAFE = false>
__device__ __forceinline__ int agent_core(const Params &p, Agent<DYN, T> &ag, const Cache<T> &c, double *tbl,
                                          uint32_t *sid, int a,
                                          const double (&obs)[Dyn<DYN>::OBS], double reward, bool act, bool learn,
                                          bool test, const Draws &d, int force, uint32_t *hash_out, uint32_t *index_out, double *o_q,
                                          double &qp, double &target, double &delta, int &swept)
{
  qp = target = delta = 0; swept = 0;
  double vp[T];
  double qn = 0;
  int own = -1, ai = 0;
  if (act)
  {
    // act on the new observation first (td.cpp:63-74), then update with (s,a,r,s',a')
    double q[A];
    if (CACHED) q_values<DYN, T, A, GEO, false>(p, tbl, obs, sid, q, ag.pidx, vp, learn, hash_out, index_out, p.n, a, CacheResolve<T>{c});
    else q_values<DYN, T, A, GEO, SAFE>(p, tbl, obs, sid, q, ag.pidx, vp, learn && !SAFE, hash_out, index_out, p.n, a, NoResolve());
    own = sample_action<A>(q, d, ag.epsilon, test, p.softmax_tau);
    ai = force >= 0 ? force : own;
    if (o_q)
    {
#pragma unroll
      for (int aa = 0; aa < A; ++aa) o_q[(size_t)aa * p.n + a] = q[aa];
    }
    qn = bootstrap_value<A>(q, ai, p.target_kind, p.target_epsilon);
  }
  else if (learn && !SAFE)
  {
#pragma unroll
    for (int j = 0; j < T; ++j) vp[j] = ld_tbl(tbl + ag.pidx[j]);
  }
  if (SAFE && learn)
  {
    int32_t *claims = p.claims + (size_t)a * p.proj.memory;
    const uint32_t mem = (uint32_t)p.proj.memory;
    // p = projector_->project(prev_obs, prev_action): claims
#pragma unroll 1
    for (int j = 0; j < T; ++j) ag.pidx[j] = probe_location(claims, mem, p.prev_hash[(size_t)j * p.n + a], true, p.pressure + a);
    if (act)
    {
      // read(projector_->project(obs, action)): claims, then the sequential sum / T of linear.cpp:147-151
      double r = 0;
#pragma unroll 1
      for (int j = 0; j < T; ++j)
      {
        const uint32_t x = probe_location(claims, mem, sid[(A * T + ai * T + j) * kMaxBlock], true, p.pressure + a);
        sid[(ai * T + j) * kMaxBlock] = x;
        r += ld_tbl(tbl + x);
      }
      qn = r / T;
    }
#pragma unroll
    for (int j = 0; j < T; ++j) vp[j] = ld_tbl(tbl + ag.pidx[j]);
  }
  if (learn)
  {
    if (CACHED) sarsa_update_cached<DYN, T>(p, ag, c, tbl, reward, act, qn, vp, qp, target, delta, swept);
    else sarsa_update<DYN, T>(p, ag, tbl, a, reward, act, qn, vp, qp, target, delta, swept);
  }
  if (act)
  {
    set_prev_projection<DYN, T>(ag, sid, ai);
    ag.action = p.actions[ai];
    if (SAFE)
    {
#pragma unroll 1
      for (int j = 0; j < T; ++j) p.prev_hash[(size_t)j * p.n + a] = sid[(A * T + ai * T + j) * kMaxBlock];
    }
  }
  return own;
}

// Episode start (online_learning.cpp:160-178): observe the start state, TDAgent::start /
// FixedAgent::start (trace clear + first act).  Used by the replay kernel; k_run inlines the same sequence.
template <int DYN, int T, int A, bool CACHED, bool GEO, bool SAFE = false>
__device__ __forceinline__ void episode_start(const Params &p, Agent<DYN, T> &ag, const Cache<T> &c, double *tbl,
                                              uint32_t *sid, int a, const Draws &d, bool test, double *o_q, int *o_ai)
{
  constexpr int OBS = Dyn<DYN>::OBS;
  double obs[OBS];
  int term;
  observe<DYN>(ag.x, p.timeout, obs, term, p.end_stop_penalty);
  if (!test)
  {
    // predictor finalize -> trace clear
    if (CACHED) cache_clear<DYN, T>(p, ag, c, tbl);
    else { ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0; }
  }
  ag.total_reward = 0;
  ag.flags = (ag.flags & 16) | 1 | (test ? 2 : 0);       // also clears bit2 (start override pending)
  double qp, target, delta; int swept;
  const int ai = agent_core<DYN, T, A, CACHED, GEO, SAFE>(p, ag, c, tbl, sid, a, obs, 0., true, false, test, d, -1, nullptr, nullptr, o_q,
                                                    qp, target, delta, swept);
  if (o_ai) *o_ai = ai;
}

// The agent half of one step (TDAgent::step / end, td.cpp:63-81; FixedAgent for test episodes).
template <int DYN, int T, int A, bool CACHED, bool GEO, bool SAFE = false>
__device__ __forceinline__ int agent_half(const Params &p, Agent<DYN, T> &ag, const Cache<T> &c, double *tbl,
                                          uint32_t *sid, int a,
                                          const double (&obs)[Dyn<DYN>::OBS], double reward, int terminal, bool test,
                                          const Draws &d, int force, uint32_t *hash_out, uint32_t *index_out, double *o_q,
                                          double &qp, double &target, double &delta, int &swept)
{
  return agent_core<DYN, T, A, CACHED, GEO, SAFE>(p, ag, c, tbl, sid, a, obs, reward, terminal != 2, !test, test, d, force,
                                            hash_out, index_out, o_q, qp, target, delta, swept);
}

// Fused K1-K4: each thread owns one agent and runs n_steps iterations of the hot loop with the
// agent's state in registers; only the hashed table and the trace ring are in memory.
// 64 threads x 7 blocks per SM keeps all 65,536 agents of BASELINE configs[1] resident at once on
// 148 SMs (448 threads/SM = 14 warps/SM).  The register file is split over the SM's four sub-partitions,
// so 14 warps means 4 warps on some of them and the kernel is bounded at 128 registers per thread
// (MINB = 7).  Launches with at most 8 warps per SM (BASELINE configs[2], [3]: 32,768 agents) use the
// MINB = 4 instantiation, which may use up to 255 registers and does not spill.
template <int DYN, int T, int A, bool CACHED, int MINB, bool GEO, bool SAFE = false>
__global__ void __launch_bounds__(kMaxBlock, MINB)
k_run(Params p, int agent_begin, int agent_end, int64_t n_steps, int do_close)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS;
  __shared__ uint32_t s_idx[(SAFE ? 2 : 1) * A * T][kMaxBlock];     // slots of Q(s', .) [+ their hashes: safe > 0]
  const int slot_in_launch = agent_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (slot_in_launch >= agent_end) return;
  // (agents are independent: which thread runs which agent cannot matter; chunk_map orders the launches by lambda)
  const int a = p.chunk_map ? (int)(__ldg(p.chunk_map + (slot_in_launch >> 5)) * 32u) + (slot_in_launch & 31) : slot_in_launch;
  uint32_t *sid = &s_idx[0][threadIdx.x];
  Agent<DYN, T> ag;
  ag.load(p, a);
  double *tbl = p.table + (size_t)a * p.tstride;
  const uint64_t gid = (uint64_t)(p.first_id + a);
  Cache<T> c;
  c.bind(p, a);
  // the cache stays live between launches, agent by agent (flags bit4): values return to the tables lazily
  if (CACHED && !(ag.flags & 16)) { cache_open<DYN, T>(p, ag, c, tbl); ag.flags |= 16; }

  // One loop iteration is either an episode start (observe the start state, first action; consumes no step) or
  // one environment step + TDAgent::step/end.  Both go through the single agent_core call below.
  for (int64_t step = 0; step < n_steps;)
  {
    const bool starting = !(ag.flags & 1);
    double obs[OBS];
    double reward = 0;
    int terminal = 0;
    bool test;
    if (starting)
    {
      // online_learning.cpp:160-178
      test = p.test_interval >= 0 && p.tt[a] % (uint64_t)(p.test_interval + 1) == (uint64_t)p.test_interval;
      const uint64_t ep = p.episode_ctr[a];
      p.episode_ctr[a] = ep + 1;
      if (p.start_override && (ag.flags & 4))
      {
#pragma unroll
        for (int i = 0; i < S; ++i) ag.x[i] = p.start_override[(size_t)i * p.n + a];
      }
      else
      {
        Philox4 w = philox_for(p.seed, gid, STREAM_START, ep);
        task_start<DYN>(u53(w.w[0], w.w[1]), p.randomization, test ? 1 : 0, ag.x);
      }
      int term0;
      observe<DYN>(ag.x, p.timeout, obs, term0, p.end_stop_penalty);
      if (!test)
      {
        // predictor finalize -> trace clear
        if (CACHED) cache_clear<DYN, T>(p, ag, c, tbl);
        else { ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0; }
      }
      ag.total_reward = 0;
      ag.flags = (ag.flags & 16) | 1 | (test ? 2 : 0);       // also clears bit2 (start override pending)
    }
    else
    {
      // ModeledEnvironment::step (modeled.cpp:160-213)
      test = (ag.flags & 2) != 0;
      double next[S];
#pragma unroll
      for (int i = 0; i < S; ++i) next[i] = ag.x[i];
      rk4<DYN>(p.tau, p.isteps, Dyn<DYN>::actuate(ag.action), next);
      observe<DYN>(next, p.timeout, obs, terminal, p.end_stop_penalty);
      reward = evaluate<DYN>(ag.x, ag.action, next, p.end_stop_penalty);
      ag.total_reward += reward;
#pragma unroll
      for (int i = 0; i < S; ++i) ag.x[i] = next[i];
    }

    double qp, target, delta; int swept;
    Draws d; d.u = 1.; d.r_explore = 0; d.r_tie = 0;
    if (terminal != 2) d = next_draws<DYN, T>(p, ag, a, gid, test);
    agent_core<DYN, T, A, CACHED, GEO, SAFE>(p, ag, c, tbl, sid, a, obs, reward, terminal != 2, !starting && !test, test, d, -1,
                                       nullptr, nullptr, nullptr, qp, target, delta, swept);
    if (!starting)
    {
      if (!test) ag.dsteps++;
      if (terminal) ag.episode_end(p, a, test);
      ++step;
    }
  }
  if (CACHED && do_close) { cache_close<DYN, T>(p, ag, c, tbl); ag.flags &= ~16; }
  ag.store(p, a);
}

// Back to the plain representation (see grlcu_cached.cuh) when the host wants to look at tables or traces:
// the fused kernel leaves the trace value cache live between launches.
template <int DYN, int T>
__global__ void k_cache_close(Params p)
{
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= p.n) return;
  if (!(p.flags[a] & 16)) return;
  Agent<DYN, T> ag;
  ag.load(p, a);
  Cache<T> c;
  c.bind(p, a);
  cache_close<DYN, T>(p, ag, c, p.table + (size_t)a * p.tstride);
  ag.flags &= ~16;
  ag.store(p, a);
}

// Replay: one agent-step per agent with states, observations, rewards, draws (and optionally the
// committed action) fed from outside; emits everything observable for comparison.
template <int DYN, int T, int A, bool GEO, bool SAFE = false>
__global__ void __launch_bounds__(kMaxBlock)
k_replay(Params p, ReplayIO io)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS;
  __shared__ uint32_t s_idx[(SAFE ? 2 : 1) * A * T][kMaxBlock];
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = p.n;
  if (a >= n) return;
  uint32_t *sid = &s_idx[0][threadIdx.x];
  Agent<DYN, T> ag;
  ag.load(p, a);
  double *tbl = p.table + (size_t)a * p.tstride;
  const bool test = io.test[a] != 0;

  if (io.started[a])
  {
#pragma unroll
    for (int i = 0; i < S; ++i) ag.x[i] = io.start_state[(size_t)i * n + a];
    Draws d; d.u = io.start_u[a]; d.r_explore = io.start_r_explore[a]; d.r_tie = io.start_r_tie[a];
    int ai;
    episode_start<DYN, T, A, false, GEO, SAFE>(p, ag, Cache<T>(), tbl, sid, a, d, test, io.o_start_q, &ai);
    if (io.o_start_action_index) io.o_start_action_index[a] = ai;
    p.episode_ctr[a] += 1;
  }

  // environment half from the re-fed state and action
  double x0[S], next[S], obs[OBS];
  int terminal;
#pragma unroll
  for (int i = 0; i < S; ++i) x0[i] = next[i] = io.state[(size_t)i * n + a];
  const double act = io.action[a];
  rk4<DYN>(p.tau, p.isteps, Dyn<DYN>::actuate(act), next);
  observe<DYN>(next, p.timeout, obs, terminal, p.end_stop_penalty);
  const double r_own = evaluate<DYN>(x0, act, next, p.end_stop_penalty);
  if (io.o_next) for (int i = 0; i < S; ++i) io.o_next[(size_t)i * n + a] = next[i];
  if (io.o_obs) for (int i = 0; i < OBS; ++i) io.o_obs[(size_t)i * n + a] = obs[i];
  if (io.o_reward) io.o_reward[a] = r_own;
  if (io.o_terminal) io.o_terminal[a] = terminal;
#pragma unroll
  for (int i = 0; i < S; ++i) ag.x[i] = next[i];

  // agent half on the fed observation / reward / terminal
#pragma unroll
  for (int i = 0; i < OBS; ++i) obs[i] = io.obs[(size_t)i * n + a];
  const double reward = io.reward[a];
  terminal = io.terminal[a];
  ag.total_reward += reward;
  double qp, target, delta; int swept;
  Draws d; d.u = io.u[a]; d.r_explore = io.r_explore[a]; d.r_tie = io.r_tie[a];
  const int force = (io.force_action && terminal != 2) ? io.force_action[a] : -1;
  const int own = agent_half<DYN, T, A, false, GEO, SAFE>(p, ag, Cache<T>(), tbl, sid, a, obs, reward, terminal, test, d, force,
                                             io.o_hash, io.o_index, io.o_q, qp, target, delta, swept);
  if (terminal != 2)
  {
    if (test) p.test_act_ctr[a] += 1; else ag.act_ctr++;
    // bounds check of every physical slot this step used (compute-sanitizer is not available on the pool)
    for (int k = 0; k < A * T; ++k) if (sid[k * kMaxBlock] >= p.tstride) *p.oob = 1;
  }
  for (int j = 0; j < T; ++j) if (ag.pidx[j] >= p.tstride) *p.oob = 1;
  if (io.o_action_index) io.o_action_index[a] = own;
  if (io.o_q_prev) io.o_q_prev[a] = qp;
  if (io.o_target) io.o_target[a] = target;
  if (io.o_delta) io.o_delta[a] = delta;
  if (io.o_swept) io.o_swept[a] = swept;
  if (io.o_trace_len) io.o_trace_len[a] = ag.len;
  if (io.o_trace_index)
    for (int k = 0; k < p.trace_cap; ++k)
    {
      const int pos = ring_pos(ag.head, k, p.trace_cap);
      for (int j = 0; j < T; ++j)
      {
        // the ring holds physical slots; the reference's trace holds the logical indices
        uint32_t x = k < ag.len ? warp_rows(p.trace_idx, a, (size_t)p.trace_cap * T)[(pos * T + j) * kLanes] : kInvalidIndex;
        if (GEO && x != kInvalidIndex) x = __ldg(p.geo.iperm + x);
        io.o_trace_index[((size_t)k * T + j) * n + a] = x;
      }
    }
  if (!test) ag.dsteps++;
  if (terminal) ag.episode_end(p, a, test);
  ag.store(p, a);
}

}  // namespace grlcu
