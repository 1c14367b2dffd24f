// grlcu_kernels.cuh -- __global__ kernels of the batched GRL learning loop (sm_100a).
#pragma once
#include "grlcu_agent.cuh"

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
__global__ void k_model_step(int64_t n, double tau, int isteps, double timeout, const double *state,
                             const double *action, double *next, double *obs, double *reward, int32_t *terminal)
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
    observe<DYN>(x, timeout, o, term);
    double r = evaluate<DYN>(x0, act, x);
#pragma unroll
    for (int i = 0; i < S; ++i) next[t * S + i] = x[i];
#pragma unroll
    for (int i = 0; i < OBS; ++i) obs[t * OBS + i] = o[i];
    reward[t] = r;
    terminal[t] = term;
  }
}

// Episode start (online_learning.cpp:160-178): Task::start, observe, TDAgent::start /
// FixedAgent::start (first act), trace clear.
template <int DYN, int T, int A>
__device__ __forceinline__ void episode_start(const Params &p, Agent<DYN, T> &ag, double *tbl, int a,
                                              const Draws &d, bool test, double *o_q, int *o_ai)
{
  constexpr int OBS = Dyn<DYN>::OBS;
  double obs[OBS], q[A];
  uint32_t idx[A][T];
  int term;
  observe<DYN>(ag.x, p.timeout, obs, term);
  if (!test) { ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0; }   // predictor finalize -> trace clear
  q_values<DYN, T, A>(p, tbl, obs, idx, q, nullptr, p.n, a);
  int ai = sample_action<A>(q, d, ag.epsilon, test);
  ag.action = p.actions[ai];
  ag.total_reward = 0;
  ag.flags = 1 | (test ? 2 : 0);       // also clears bit2 (start override pending)
  set_prev_projection<DYN, T, A>(ag, idx, ai);
  if (o_q)
  {
#pragma unroll
    for (int aa = 0; aa < A; ++aa) o_q[(size_t)aa * p.n + a] = q[aa];
  }
  if (o_ai) *o_ai = ai;
}

__device__ __forceinline__ Draws draws_from(const Philox4 &w)
{
  Draws d;
  d.u = u53(w.w[0], w.w[1]);
  d.r_explore = w.w[2] >> 1;
  d.r_tie = w.w[3] >> 1;
  return d;
}

// Fused K1-K4: each thread owns one agent and runs n_steps iterations of the hot loop with
// the agent's state in registers; only the hashed table and the trace ring are in memory.
template <int DYN, int T, int A>
__global__ void __launch_bounds__(128)
k_run(Params p, int agent_begin, int agent_end, int64_t n_steps)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS;
  const int a = agent_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= agent_end) return;
  Agent<DYN, T> ag;
  ag.load(p, a);
  double *tbl = p.table + (size_t)a * p.proj.memory;
  const uint64_t gid = (uint64_t)(p.first_id + a);

  for (int64_t step = 0; step < n_steps; ++step)
  {
    if (!(ag.flags & 1))
    {
      const bool test = p.test_interval >= 0 &&
                        ag.tt % (uint64_t)(p.test_interval + 1) == (uint64_t)p.test_interval;
      if (p.start_override && (ag.flags & 4))
      {
#pragma unroll
        for (int i = 0; i < S; ++i) ag.x[i] = p.start_override[(size_t)i * p.n + a];
      }
      else
      {
        Philox4 w = philox_for(p.seed, gid, STREAM_START, ag.episode_ctr);
        task_start<DYN>(u53(w.w[0], w.w[1]), p.randomization, test ? 1 : 0, ag.x);
      }
      ag.episode_ctr++;
      Philox4 w = test ? philox_for(p.seed, gid, STREAM_TEST_ACT, ag.test_act_ctr++)
                       : philox_for(p.seed, gid, STREAM_ACT, ag.act_ctr++);
      episode_start<DYN, T, A>(p, ag, tbl, a, draws_from(w), test, nullptr, nullptr);
    }
    const bool test = (ag.flags & 2) != 0;

    // ModeledEnvironment::step (modeled.cpp:160-213)
    double next[S], obs[OBS];
    int terminal;
#pragma unroll
    for (int i = 0; i < S; ++i) next[i] = ag.x[i];
    rk4<DYN>(p.tau, p.isteps, Dyn<DYN>::actuate(ag.action), next);
    observe<DYN>(next, p.timeout, obs, terminal);
    const double reward = evaluate<DYN>(ag.x, ag.action, next);
    ag.total_reward += reward;

    double qp, target, delta; int swept;
    if (terminal == 2)
    {
      // TDAgent::end (td.cpp:76-81)
      if (!test) sarsa_update<DYN, T>(p, ag, tbl, a, reward, false, 0., qp, target, delta, swept);
    }
    else
    {
      // TDAgent::step (td.cpp:63-74): act, then update
      double q[A];
      uint32_t idx[A][T];
      q_values<DYN, T, A>(p, tbl, obs, idx, q, nullptr, p.n, a);
      Philox4 w = test ? philox_for(p.seed, gid, STREAM_TEST_ACT, ag.test_act_ctr++)
                       : philox_for(p.seed, gid, STREAM_ACT, ag.act_ctr++);
      const int ai = sample_action<A>(q, draws_from(w), ag.epsilon, test);
      double qn = q[0];
#pragma unroll
      for (int aa = 1; aa < A; ++aa) if (aa == ai) qn = q[aa];
      if (!test) sarsa_update<DYN, T>(p, ag, tbl, a, reward, true, qn, qp, target, delta, swept);
      set_prev_projection<DYN, T, A>(ag, idx, ai);
      ag.action = p.actions[ai];
    }
#pragma unroll
    for (int i = 0; i < S; ++i) ag.x[i] = next[i];
    if (!test) ag.ss++;
    if (terminal)
    {
      ag.flags = 0;
      ag.tt++;
      ag.episodes++;
      if (test) ag.last_test_return = ag.total_reward;
      else { ag.last_return = ag.total_reward; ag.return_sum += ag.total_reward; }
    }
  }
  ag.store(p, a);
}

// Replay: one agent-step per agent with states, observations, rewards, draws (and optionally the
// committed action) fed from outside; emits everything observable for comparison.
template <int DYN, int T, int A>
__global__ void __launch_bounds__(128)
k_replay(Params p, ReplayIO io)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS;
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = p.n;
  if (a >= n) return;
  Agent<DYN, T> ag;
  ag.load(p, a);
  double *tbl = p.table + (size_t)a * p.proj.memory;
  const bool test = io.test[a] != 0;

  if (io.started[a])
  {
#pragma unroll
    for (int i = 0; i < S; ++i) ag.x[i] = io.start_state[(size_t)i * n + a];
    Draws d; d.u = io.start_u[a]; d.r_explore = io.start_r_explore[a]; d.r_tie = io.start_r_tie[a];
    int ai;
    episode_start<DYN, T, A>(p, ag, tbl, a, d, test, io.o_start_q, &ai);
    if (io.o_start_action_index) io.o_start_action_index[a] = ai;
    ag.episode_ctr++;
  }

  // environment half from the re-fed state and action
  double x0[S], next[S], obs[OBS];
  int terminal;
#pragma unroll
  for (int i = 0; i < S; ++i) x0[i] = next[i] = io.state[(size_t)i * n + a];
  const double act = io.action[a];
  rk4<DYN>(p.tau, p.isteps, Dyn<DYN>::actuate(act), next);
  observe<DYN>(next, p.timeout, obs, terminal);
  const double r_own = evaluate<DYN>(x0, act, next);
  if (io.o_next) for (int i = 0; i < S; ++i) io.o_next[(size_t)i * n + a] = next[i];
  if (io.o_obs) for (int i = 0; i < OBS; ++i) io.o_obs[(size_t)i * n + a] = obs[i];
  if (io.o_reward) io.o_reward[a] = r_own;
  if (io.o_terminal) io.o_terminal[a] = terminal;

  // agent half on the fed observation / reward / terminal
#pragma unroll
  for (int i = 0; i < OBS; ++i) obs[i] = io.obs[(size_t)i * n + a];
  const double reward = io.reward[a];
  terminal = io.terminal[a];
  ag.total_reward += reward;
  double qp = 0, target = 0, delta = 0; int swept = 0, own = -1;
  if (terminal == 2)
  {
    if (!test) sarsa_update<DYN, T>(p, ag, tbl, a, reward, false, 0., qp, target, delta, swept);
  }
  else
  {
    double q[A];
    uint32_t idx[A][T];
    q_values<DYN, T, A>(p, tbl, obs, idx, q, io.o_hash, n, a);
    Draws d; d.u = io.u[a]; d.r_explore = io.r_explore[a]; d.r_tie = io.r_tie[a];
    own = sample_action<A>(q, d, ag.epsilon, test);
    const int ai = io.force_action ? io.force_action[a] : own;
    for (int aa = 0; aa < A; ++aa)
    {
      if (io.o_q) io.o_q[(size_t)aa * n + a] = q[aa];
      if (io.o_index) for (int j = 0; j < T; ++j) io.o_index[((size_t)aa * T + j) * n + a] = idx[aa][j];
    }
    double qn = q[0];
#pragma unroll
    for (int aa = 1; aa < A; ++aa) if (aa == ai) qn = q[aa];
    if (!test) sarsa_update<DYN, T>(p, ag, tbl, a, reward, true, qn, qp, target, delta, swept);
    set_prev_projection<DYN, T, A>(ag, idx, ai);
    ag.action = p.actions[ai];
    if (test) ag.test_act_ctr++; else ag.act_ctr++;
  }
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
        io.o_trace_index[((size_t)k * T + j) * n + a] =
            k < ag.len ? p.trace_idx[((size_t)pos * T + j) * n + a] : kInvalidIndex;
    }
#pragma unroll
  for (int i = 0; i < S; ++i) ag.x[i] = next[i];
  if (!test) ag.ss++;
  if (terminal)
  {
    ag.flags = 0; ag.tt++; ag.episodes++;
    if (test) ag.last_test_return = ag.total_reward;
    else { ag.last_return = ag.total_reward; ag.return_sum += ag.total_reward; }
  }
  ag.store(p, a);
}

}  // namespace grlcu
