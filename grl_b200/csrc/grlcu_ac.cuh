// grlcu_ac.cuh -- actor-critic variant of the agent-step (BASELINE configs[3]: cart-double-pole AC with two
// tile-coded tables per agent).  Reference: agent/td (agents/td.cpp:50-81) around
//   policy/parameterized/action   ActionPolicy::act  (policies/action.cpp:99-158): a = clamp(pi(s) + n),
//                                 n = (1-theta) n + N(0, sigma) by Box-Muller (utils.h:120-125), theta = 1
//   predictor/ac/action           ActionACPredictor::update (predictors/ac.cpp:72-110), proportional:
//                                 u = pi(s); delta = critic(...); write(ap, u + delta (a - u), alpha_actor)
//   predictor/critic/td           TDPredictor::criticize (predictors/td.cpp:68-91): TD(lambda) on V(s) with
//                                 the same write / trace sweep / trace add as SARSA
// Actor and critic use projectors with the same parameters (cfg/cart_double_pole/ac.yaml:47-53 derives the
// critic's from the actor's), and TileCodingProjector hashes without a per-object seed, so one set of T
// indices addresses both per-agent tables.  The critic update is sarsa_update() (grlcu_agent.cuh) verbatim.
// The actor table is clamped to [action_min, action_max] on every write (linear.cpp:207-211, limit = 1)
// and on every read (linear.cpp:179-180).
//
// Physical layout: because one index x addresses both tables and the agent always reads (and writes) the critic's and the
// actor's word of an index together, the two tables of an agent are INTERLEAVED: pair x = {critic[x], actor[x]} at doubles
// [2x, 2x + 1] of one array (`critic` = base, `actor` = base + 1, physical slot = 2x).  One 16-byte request fetches both
// (ld_tbl2), and the actor store and the critic's directory sentinel of an index dirty ONE sector.  The logical index x
// (replay, grlcu_get_table / get_actor_table, checkpoints) is slot >> 1; the trace ring and prev_idx hold slots.
#pragma once
#include "grlcu_agent.cuh"
#include "grlcu_cached.cuh"

namespace grlcu {

struct ACParams
{
  double *actor;                 // [n][memory]
  double  sigma, actor_alpha, action_min, action_max;
};

// TileCodingProjector::_project of an observation (no action coordinate): OBS + 1 ints per hash.  Rolled over
// the tilings (code size, see q_values); the indices go to the shared-memory stash sid[j * kMaxBlock].
template <int OBS, int T>
__device__ __forceinline__ void project_obs(const Params &p, const double (&obs)[OBS], uint32_t *sid,
                                            uint32_t *hash_out, size_t stride)
{
  int32_t q[OBS];
  quantize<OBS, T>(p.proj, obs, q);
#pragma unroll 1
  for (int j = 0; j < T; ++j)
  {
    const uint32_t h = tile_hash_of<OBS, T>(p.proj, q, j);
    if (hash_out) hash_out[(size_t)j * stride] = h;
    sid[j * kMaxBlock] = reduce_index(p, h) * 2u;       // physical slot of the {critic, actor} pair
  }
}

__device__ __forceinline__ double clampd(double v, double lo, double hi) { return fmin(fmax(v, lo), hi); }

// Rand::getNormal(0, sigma) from two uniforms (utils.h:120-125)
__device__ __forceinline__ double box_muller(double U1, double U2, double sigma)
{
  return sqrt(-2 * log(U1)) * cos(2 * GRLCU_PI * U2) * sigma + 0.;
}

struct ACOut { double pi, vnext, u, action; };

// ActionACPredictor::update for the transition (p, prev_action, reward [, s', a']).
//   ua[j] / vp[j]: actor / critic table words at pidx[j], read before any store of this step (vp raw: value or
//   directory sentinel when CACHED).
template <int DYN, int T, bool CACHED>
__device__ __forceinline__ void ac_update(const Params &p, const ACParams &ac, Agent<DYN, T> &ag, const Cache<T> &c,
                                          double *critic, double *actor, int a, double reward, bool bootstrap,
                                          double vnext, const double (&ua)[T], const double (&vp)[T],
                                          double &u_out, double &qp, double &target, double &delta, int &swept)
{
  // u = representation_->read(ap): sequential sum / T, clamped
  double u = 0;
#pragma unroll
  for (int j = 0; j < T; ++j) u += ua[j];
  u = clampd(u / T, ac.action_min, ac.action_max);
  // delta = critique * (prev_action - u); target_u = u + delta; representation_->write(ap, target_u, alpha):
  // the reference does this after critic_->criticize() returns; the actor table is not touched by the critic, so
  // the stores are issued as soon as the critique is known (cached path: before the critic's trace sweep, while the
  // sectors fetched with ua are still in L2)
  auto actor_write = [&](double critique)
  {
    const double du = critique * (ag.action - u);
    const double target_u = u + du;
    const double dw = ac.actor_alpha * (target_u - u);    // write() reads the same, unchanged, actor value again
    if (!(ag.flags & 8))
    {
#pragma unroll
      for (int j = 0; j < T; ++j) actor[ag.pidx[j]] = clampd(ua[j] + dw, ac.action_min, ac.action_max);
    }
    else
    {
      uint32_t lp[T];
#pragma unroll
      for (int j = 0; j < T; ++j) lp[j] = ag.pidx[j];
#pragma unroll 1
      for (int j = 0; j < T; ++j)                          // duplicate indices: sequential, clamped each time
        actor[lp[j]] = clampd(ld_tbl(actor + lp[j]) + dw, ac.action_min, ac.action_max);
    }
  };
  // critique = critic_->criticize(transition, u)
  if (CACHED) sarsa_update_cached<DYN, T>(p, ag, c, critic, reward, bootstrap, vnext, vp, qp, target, delta, swept, actor_write);
  else
  {
    sarsa_update<DYN, T>(p, ag, critic, a, reward, bootstrap, vnext, vp, qp, target, delta, swept);
    actor_write(delta);
  }
  ag.dtblw += (uint32_t)T;                                  // actor stores (the critic's are counted by sarsa_update)
  u_out = u;
}

// The agent's part of one loop iteration, written once (see agent_core in grlcu_kernels.cuh):
//   act   : ActionPolicy::act on obs (learning: with exploration noise n; test: sigma = 0)
//   learn : the update that TDAgent::step / end performs right after it
// have_n: n is supplied (replay).
template <int DYN, int T, bool CACHED>
__device__ __forceinline__ void ac_core(const Params &p, const ACParams &ac, Agent<DYN, T> &ag, const Cache<T> &c,
                                        double *critic, double *actor, uint32_t *sid, int a, uint64_t gid,
                                        const double (&obs)[Dyn<DYN>::OBS], double reward, bool act, bool learn,
                                        bool test, bool have_n, double n_fed,
                                        uint32_t *hash_out, uint32_t *index_out, ACOut &o, double &noise,
                                        double &qp, double &target, double &delta, int &swept)
{
  constexpr int OBS = Dyn<DYN>::OBS;
  qp = target = delta = 0; swept = 0; noise = 0;
  o.pi = o.vnext = o.u = o.action = 0;
  double pi = 0, vnext = 0, action = 0;
  if (act)
  {
    project_obs<OBS, T>(p, obs, sid, hash_out, (size_t)p.n);
    double wa[T], wc[T];
#pragma unroll
    for (int j = 0; j < T; ++j)
    {
      const double2 w = ld_tbl2(critic + sid[j * kMaxBlock]);     // {critic, actor} of one index: one request
      wc[j] = w.x; wa[j] = w.y;
    }
#pragma unroll
    for (int j = 0; j < T; ++j) { pi += wa[j]; vnext += CACHED ? resolve<T>(c, wc[j]) : wc[j]; }
    pi = clampd(pi / T, ac.action_min, ac.action_max);
    vnext /= T;
    double out = pi;
    if (!test && ac.sigma != 0.)
    {
      double n;
      if (have_n) n = n_fed;
      else
      {
        const Philox4 w = philox_for(p.seed, gid, STREAM_NOISE, ag.act_ctr);
        n = (1 - 1.) * 0. + box_muller(u53(w.w[0], w.w[1]), u53(w.w[2], w.w[3]), 1. * ac.sigma);
      }
      ag.act_ctr++;
      out += n;
      noise = n;
    }
    action = clampd(out, ac.action_min, ac.action_max);
  }
  if (learn)
  {
    double ua[T], vp[T];
#pragma unroll
    for (int j = 0; j < T; ++j) { const double2 w = ld_tbl2(critic + ag.pidx[j]); vp[j] = w.x; ua[j] = w.y; }
    ac_update<DYN, T, CACHED>(p, ac, ag, c, critic, actor, a, reward, act, vnext, ua, vp, o.u, qp, target, delta, swept);
  }
  if (act)
  {
#pragma unroll
    for (int j = 0; j < T; ++j) ag.pidx[j] = sid[j * kMaxBlock];
    ag.flags = (ag.flags & ~8) | (has_duplicates<T>(ag.pidx) ? 8 : 0);
    ag.action = action;
    if (index_out)
    {
#pragma unroll
      for (int j = 0; j < T; ++j) index_out[(size_t)j * p.n] = ag.pidx[j] >> 1;     // logical index
    }
    o.pi = pi; o.vnext = vnext; o.action = action;
  }
}

// Fused loop, one thread per agent: n_steps iterations of online_learning.cpp:186-221 with the AC agent.
// One loop iteration is an episode start (no step consumed) or one environment step + TDAgent::step/end.
template <int DYN, int T, bool CACHED, int MINB>
__global__ void __launch_bounds__(kMaxBlock, MINB)
k_run_ac(Params p, ACParams ac, int agent_begin, int agent_end, int64_t n_steps, int do_close)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS;
  __shared__ uint32_t s_idx[T][kMaxBlock];
  const int a = agent_begin + blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= agent_end) return;
  uint32_t *sid = &s_idx[0][threadIdx.x];
  Agent<DYN, T> ag;
  ag.load(p, a);
  double *critic = p.table + (size_t)a * p.tstride;
  double *actor = critic + 1;                  // interleaved pairs
  const uint64_t gid = (uint64_t)(p.first_id + a);
  Cache<T> c;
  c.bind(p, a);
  if (CACHED && !(ag.flags & 16)) { cache_open<DYN, T>(p, ag, c, critic); ag.flags |= 16; }
  ACOut o;
  double noise, qp, target, delta;
  int swept;

  for (int64_t step = 0; step < n_steps;)
  {
    const bool starting = !(ag.flags & 1);
    double obs[OBS];
    double reward = 0;
    int terminal = 0;
    bool test;
    if (starting)
    {
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
      // TDAgent::start -> predictor_->finalize() is ActionACPredictor::finalize() (ac.cpp:170-173): the empty
      // Predictor::finalize() and nothing else -- it does not reach the critic, so the critic's eligibility trace is NOT
      // cleared between episodes (TDPredictor::finalize, td.cpp:93-99, only runs on the experiment's reset at the start of
      // a run = grlcu_reset).  The reference's own ac.cpp / td.cpp, compiled, say so (tests/test_oracle_vs_ref_agent.py).
      ag.total_reward = 0;
      ag.flags = (ag.flags & 16) | 1 | (test ? 2 : 0);
    }
    else
    {
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
    ac_core<DYN, T, CACHED>(p, ac, ag, c, critic, actor, sid, a, gid, obs, reward, terminal != 2, !starting && !test,
                            test, false, 0., nullptr, nullptr, o, noise, qp, target, delta, swept);
    if (!starting)
    {
      if (!test) ag.dsteps++;
      if (terminal) ag.episode_end(p, a, test);
      ++step;
    }
  }
  if (CACHED && do_close) { cache_close<DYN, T>(p, ag, c, critic); ag.flags &= ~16; }
  ag.store(p, a);
}

// Replay: one agent-step per agent with state, action, observation, reward, terminal and the exploration
// noise fed from outside.  ReplayIO use in AC mode: start_u / u carry the noise n; o_hash / o_index hold
// one row of T (the projection of the fed observation); o_q rows: pi(s'), V(s'), u, a'; o_start_q rows:
// pi(s0), -, -, a0.
template <int DYN, int T>
__global__ void __launch_bounds__(kMaxBlock)
k_replay_ac(Params p, ACParams ac, ReplayIO io)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS;
  __shared__ uint32_t s_idx[T][kMaxBlock];
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = p.n;
  if (a >= n) return;
  uint32_t *sid = &s_idx[0][threadIdx.x];
  Agent<DYN, T> ag;
  ag.load(p, a);
  double *critic = p.table + (size_t)a * p.tstride;
  double *actor = critic + 1;                  // interleaved pairs
  const uint64_t gid = (uint64_t)(p.first_id + a);
  const bool test = io.test[a] != 0;
  ACOut o;
  double noise, qp, target, delta;
  int swept;

  if (io.started[a])
  {
#pragma unroll
    for (int i = 0; i < S; ++i) ag.x[i] = io.start_state[(size_t)i * n + a];
    double obs0[OBS];
    int term0;
    observe<DYN>(ag.x, p.timeout, obs0, term0, p.end_stop_penalty);
    // (no trace clear: ActionACPredictor::finalize does not reach the critic, see k_run_ac)
    ag.total_reward = 0;
    ag.flags = (ag.flags & 16) | 1 | (test ? 2 : 0);
    ac_core<DYN, T, false>(p, ac, ag, Cache<T>(), critic, actor, sid, a, gid, obs0, 0., true, false, test, true,
                           io.start_u[a], nullptr, nullptr, o, noise, qp, target, delta, swept);
    if (io.o_start_q) { io.o_start_q[a] = o.pi; io.o_start_q[(size_t)3 * n + a] = o.action; }
    p.episode_ctr[a] += 1;
  }

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

#pragma unroll
  for (int i = 0; i < OBS; ++i) obs[i] = io.obs[(size_t)i * n + a];
  const double reward = io.reward[a];
  terminal = io.terminal[a];
  ag.total_reward += reward;
  ag.action = act;                       // the action that was applied is the transition's prev_action
  ac_core<DYN, T, false>(p, ac, ag, Cache<T>(), critic, actor, sid, a, gid, obs, reward, terminal != 2, !test, test,
                         true, io.u[a], io.o_hash ? io.o_hash + a : nullptr, io.o_index ? io.o_index + a : nullptr, o,
                         noise, qp, target, delta, swept);
  for (int j = 0; j < T; ++j) if (ag.pidx[j] + 1u >= p.tstride + 0u && terminal != 2) *p.oob = 1;     // pair {critic, actor} inside the row
  if (io.o_q)
  {
    io.o_q[a] = o.pi; io.o_q[(size_t)n + a] = o.vnext; io.o_q[(size_t)2 * n + a] = o.u; io.o_q[(size_t)3 * n + a] = o.action;
  }
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
        uint32_t x = k < ag.len ? warp_rows(p.trace_idx, a, (size_t)p.trace_cap * T)[(pos * T + j) * kLanes] : kInvalidIndex;
        if (x != kInvalidIndex) x >>= 1;           // the ring holds physical slots
        io.o_trace_index[((size_t)k * T + j) * n + a] = x;
      }
    }
  if (!test) ag.dsteps++;
  if (terminal) ag.episode_end(p, a, test);
  ag.store(p, a);
}

}  // namespace grlcu
