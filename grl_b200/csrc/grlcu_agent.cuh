// grlcu_agent.cuh -- one agent-step (online_learning.cpp:186-221) for one agent per thread.
//
// Layout in HBM (all per-context, structure-of-arrays over agents so that a warp's 32
// agents read/write consecutive addresses for everything except the hashed tables):
//   table      double  [n][memory]          per-agent hashed weight table (agent-major)
//   state      double  [S][n]               environment state
//   action     double  [n]                  action to apply at the next step
//   prev_idx   uint32  [T][n]               projection of (prev_obs, prev_action)
//   trace_idx  uint32  [cap][T][n]          enumerated trace ring (0xFFFFFFFF = invalidated)
//   trace_len/head int [n], total_decay double [n]
//   counters / returns [n]
// Citations: /root/reference/base.
#pragma once
#include "grlcu_device.cuh"

namespace grlcu {

struct Params
{
  int32_t  n;                 // agents in this context
  int64_t  first_id;          // global id of agent 0 (Philox stream)
  uint64_t seed;
  double   tau; int32_t isteps; double timeout, randomization;
  Projector proj;
  double   actions[8];
  int32_t  test_interval;
  double   alpha, gamma, lambda, epsilon;
  int32_t  trace_kind, trace_cap;
  // per-agent state
  double   *table;
  double   *state, *action;
  uint32_t *prev_idx, *trace_idx;
  int32_t  *trace_len, *trace_head;
  double   *total_decay;
  int32_t  *flags;            // bit0 in_episode, bit1 test, bit2 start override pending, bit3 prev projection has duplicate indices
  uint32_t *dupmask;          // bit k: trace entry of age k holds duplicate indices
  int32_t  force_slow;        // 1: always take the sequential update path (A/B testing)
  uint64_t *act_ctr, *test_act_ctr, *episode_ctr, *tt, *ss, *swept;
  uint32_t *episodes;
  double   *total_reward, *last_return, *last_test_return, *return_sum;
  const double *alpha_a, *lambda_a, *epsilon_a;   // optional per-agent hyper-parameters
  const double *start_override;                    // optional [S][n] start states (run_host)
};

// Replay buffers on the device (same meaning as grlcu_replay_in/out, SoA [field][agent])
struct ReplayIO
{
  const int32_t *started; const double *start_state; const double *start_u;
  const uint32_t *start_r_explore, *start_r_tie;
  const double *state, *action; const int32_t *test;
  const double *obs, *reward; const int32_t *terminal;
  const double *u; const uint32_t *r_explore, *r_tie; const int32_t *force_action;
  double *o_next, *o_obs, *o_reward; int32_t *o_terminal;
  uint32_t *o_hash, *o_index; double *o_q; int32_t *o_action_index;
  double *o_start_q; int32_t *o_start_action_index;
  double *o_q_prev, *o_target, *o_delta; int32_t *o_swept, *o_trace_len; uint32_t *o_trace_index;
};

struct Draws { double u; uint32_t r_explore, r_tie; };

// greedy.cpp:47-61 findmax + 63-86 / 151-218 sample (epsilon_.size() == 1, decay_ = 1)
template <int A>
__device__ __forceinline__ int sample_action(const double (&q)[A], const Draws &d, double epsilon, bool test)
{
  int mai = 0, man = 1;
#pragma unroll
  for (int ii = 1; ii < A; ++ii)
  {
    if (q[ii] > q[mai]) { mai = ii; man = 1; }
    else if (q[ii] == q[mai]) man++;
  }
  if (!test && d.u < 1. * epsilon)
    return (int)(d.r_explore % (uint32_t)A);
  if (man > 1)
  {
    int jj = (int)(d.r_tie % (uint32_t)man);
    int ii = mai;
    for (; jj >= 0; ++ii)
      if (q[ii] == q[mai]) --jj;
    return ii - 1;
  }
  return mai;
}

// Per-thread view of one agent's persistent state.
template <int DYN, int T>
struct Agent
{
  static constexpr int S = Dyn<DYN>::S;
  double   x[S];
  double   action;
  uint32_t pidx[T];
  int32_t  len, head, flags;
  uint32_t dupmask;
  double   total_decay, total_reward;
  uint64_t act_ctr, test_act_ctr, episode_ctr, tt, ss, swept;
  uint32_t episodes;
  double   last_return, last_test_return, return_sum;
  double   alpha, lambda, epsilon;

  __device__ __forceinline__ void load(const Params &p, int a)
  {
#pragma unroll
    for (int i = 0; i < S; ++i) x[i] = p.state[(size_t)i * p.n + a];
    action = p.action[a];
#pragma unroll
    for (int j = 0; j < T; ++j) pidx[j] = p.prev_idx[(size_t)j * p.n + a];
    len = p.trace_len[a]; head = p.trace_head[a]; flags = p.flags[a]; dupmask = p.dupmask[a];
    total_decay = p.total_decay[a]; total_reward = p.total_reward[a];
    act_ctr = p.act_ctr[a]; test_act_ctr = p.test_act_ctr[a]; episode_ctr = p.episode_ctr[a];
    tt = p.tt[a]; ss = p.ss[a]; swept = p.swept[a]; episodes = p.episodes[a];
    last_return = p.last_return[a]; last_test_return = p.last_test_return[a]; return_sum = p.return_sum[a];
    alpha = p.alpha_a ? p.alpha_a[a] : p.alpha;
    lambda = p.lambda_a ? p.lambda_a[a] : p.lambda;
    epsilon = p.epsilon_a ? p.epsilon_a[a] : p.epsilon;
  }
  __device__ __forceinline__ void store(const Params &p, int a) const
  {
#pragma unroll
    for (int i = 0; i < S; ++i) p.state[(size_t)i * p.n + a] = x[i];
    p.action[a] = action;
#pragma unroll
    for (int j = 0; j < T; ++j) p.prev_idx[(size_t)j * p.n + a] = pidx[j];
    p.trace_len[a] = len; p.trace_head[a] = head; p.flags[a] = flags; p.dupmask[a] = dupmask;
    p.total_decay[a] = total_decay; p.total_reward[a] = total_reward;
    p.act_ctr[a] = act_ctr; p.test_act_ctr[a] = test_act_ctr; p.episode_ctr[a] = episode_ctr;
    p.tt[a] = tt; p.ss[a] = ss; p.swept[a] = swept; p.episodes[a] = episodes;
    p.last_return[a] = last_return; p.last_test_return[a] = last_test_return; p.return_sum[a] = return_sum;
  }
};

// QPolicy::values (q.cpp:94-107): projections of (obs, a) for every action, then
// LinearRepresentation::read (linear.cpp:136-184): sequential sum in tiling order, / T.
// All A*T gathers are issued before the first sum so they are in flight together.
template <int DYN, int T, int A>
__device__ __forceinline__ void q_values(const Params &p, const double *tbl,
                                         const double (&obs)[Dyn<DYN>::OBS],
                                         uint32_t (&idx)[A][T], double (&q)[A], uint32_t *hash_out, int n, int a)
{
  constexpr int OBS = Dyn<DYN>::OBS, DIMS = OBS + 1;
  double w[A][T];
#pragma unroll
  for (int aa = 0; aa < A; ++aa)
  {
    double in[DIMS];
    int32_t qz[DIMS];
#pragma unroll
    for (int i = 0; i < OBS; ++i) in[i] = obs[i];
    in[OBS] = p.actions[aa];
    quantize<DIMS, T>(p.proj, in, qz);
#pragma unroll
    for (int j = 0; j < T; ++j)
    {
      uint32_t h = tile_hash_of<DIMS, T>(p.proj, qz, j);
      idx[aa][j] = h % (uint32_t)p.proj.memory;
      if (hash_out) hash_out[((size_t)aa * T + j) * n + a] = h;
      w[aa][j] = tbl[idx[aa][j]];
    }
  }
#pragma unroll
  for (int aa = 0; aa < A; ++aa)
  {
    double r = 0;
#pragma unroll
    for (int j = 0; j < T; ++j) r += w[aa][j];
    q[aa] = r / T;      // clamp to +-DBL_MAX is the identity for finite values (linear.cpp:179-180)
  }
}

__device__ __forceinline__ int ring_pos(int head, int k, int cap)
{
  int pos = head - k;
  return pos < 0 ? pos + cap : pos;
}

// true if any two of the T indices are equal (hash collision between tilings)
template <int T>
__device__ __forceinline__ bool has_duplicates(const uint32_t (&ix)[T])
{
  bool dup = false;
#pragma unroll
  for (int i = 1; i < T; ++i)
#pragma unroll
    for (int j = 0; j < i; ++j) dup |= (ix[i] == ix[j]);
  return dup;
}

// Install the projection chosen by the policy as the agent's (prev_obs, prev_action) projection.
template <int DYN, int T, int A>
__device__ __forceinline__ void set_prev_projection(Agent<DYN, T> &ag, const uint32_t (&idx)[A][T], int ai)
{
#pragma unroll
  for (int aa = 0; aa < A; ++aa)
    if (aa == ai)
    {
#pragma unroll
      for (int j = 0; j < T; ++j) ag.pidx[j] = idx[aa][j];
    }
  ag.flags = (ag.flags & ~8) | (has_duplicates<T>(ag.pidx) ? 8 : 0);
}

// EnumeratedTrace::add tail (trace.h:226-233 / 253-260): push p, total_decay *= decay, pop while
// total_decay < dc.  The ring only moves its head; popped entries are simply no longer counted.
template <int DYN, int T>
__device__ __forceinline__ void trace_push(const Params &p, Agent<DYN, T> &ag, uint32_t *tr, double e, double dc)
{
  const int n = p.n, cap = p.trace_cap;
  ag.head = ag.head + 1 == cap ? 0 : ag.head + 1;
#pragma unroll
  for (int j = 0; j < T; ++j) tr[((size_t)ag.head * T + j) * n] = ag.pidx[j];
  ag.len++;
  ag.dupmask = (ag.dupmask << 1) | ((ag.flags & 8) ? 1u : 0u);
  ag.total_decay *= e;
  while (ag.total_decay < dc && ag.len > 1) { ag.total_decay /= e; ag.len--; }
}

// SARSAPredictor::criticize (sarsa.cpp:98-124) with tau = 1:
//   target = r [+ gamma*Q(s',a')]; delta = target - Q(p); write(p, target, alpha);
//   update(trace, alpha*delta, gamma*lambda); trace.add(p, gamma*lambda)
//
// The reference applies every table increment as a sequential read-modify-write (p's T entries, then
// the trace entries newest to oldest).  Two code paths give bit-identical tables:
//  * slow path: exactly that sequence (needed whenever one address can receive two increments whose
//    order matters: duplicate indices inside a projection, or the accumulating trace);
//  * fast path (replacing trace, no duplicate indices in p or in any live entry): all valid trace
//    indices are then pairwise distinct (ssub removed every index an older entry shared with a newer
//    one, projection.h:93-104), so after p's T stores every remaining address gets exactly one
//    increment and the gathers/scatters are independent: they are issued G entries (G*T loads) at a
//    time instead of one dependent RMW after another.
template <int DYN, int T>
__device__ __forceinline__ void sarsa_update(const Params &p, Agent<DYN, T> &ag, double *tbl,
                                             int a, double reward, bool bootstrap, double q_next,
                                             double &o_qp, double &o_target, double &o_delta, int &o_swept)
{
  const int n = p.n, cap = p.trace_cap;
  uint32_t *tr = p.trace_idx + a;               // element (pos, j) at tr[((size_t)pos*T + j)*n]
  double target = reward;
  if (bootstrap) target += p.gamma * q_next;    // pow(gamma, 1)
  double vp[T];
#pragma unroll
  for (int j = 0; j < T; ++j) vp[j] = tbl[ag.pidx[j]];
  double qp = 0;
#pragma unroll
  for (int j = 0; j < T; ++j) qp += vp[j];
  qp /= T;
  const double delta = target - qp;
  // LinearRepresentation::write (linear.cpp:186-196): delta_w = alpha*(target - read(p))
  const double dw = ag.alpha * (target - qp);
  const double e = p.gamma * ag.lambda;         // pow(gamma*lambda, 1)
  const double ad = ag.alpha * delta;
  const double dc = p.trace_kind == 1 ? 0.01 : 0.0001;
  const uint32_t live = ag.len >= 32 ? 0xFFFFFFFFu : ((1u << ag.len) - 1u);
  const bool fast = !p.force_slow && p.trace_kind != 2 && cap <= 32 && !(ag.flags & 8) &&
                    (ag.dupmask & live) == 0;
  int swept = 0;

  if (fast)
  {
#pragma unroll
    for (int j = 0; j < T; ++j) tbl[ag.pidx[j]] = vp[j] + dw;
    if (p.trace_kind == 1)
    {
      constexpr int G = 2;
      double weight = 1.;
      for (int k0 = 0; k0 < ag.len; k0 += G)
      {
        uint32_t ix[G][T];
        double v[G][T], inc[G];
        bool sw[G];
#pragma unroll
        for (int g = 0; g < G; ++g)
        {
          const int k = k0 + g;
          const bool in = k < ag.len;
          const int pos = ring_pos(ag.head, in ? k : 0, cap);
          sw[g] = in && weight > 0.001;
          inc[g] = weight * ad * e;
          weight *= e;
#pragma unroll
          for (int j = 0; j < T; ++j) ix[g][j] = in ? tr[((size_t)pos * T + j) * n] : kInvalidIndex;
        }
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
          for (int j = 0; j < T; ++j)
            if (sw[g] && ix[g][j] != kInvalidIndex) v[g][j] = tbl[ix[g][j]];
#pragma unroll
        for (int g = 0; g < G; ++g)
        {
          const int k = k0 + g;
          const int pos = ring_pos(ag.head, k < ag.len ? k : 0, cap);
#pragma unroll
          for (int j = 0; j < T; ++j)
          {
            const uint32_t x = ix[g][j];
            if (x != kInvalidIndex)
            {
              if (sw[g]) { tbl[x] = v[g][j] + inc[g]; ++swept; }
              bool hit = false;                                   // ssub against p (projection.h:93-104)
#pragma unroll
              for (int jj = 0; jj < T; ++jj) hit |= (x == ag.pidx[jj]);
              if (hit) tr[((size_t)pos * T + j) * n] = kInvalidIndex;
            }
          }
        }
      }
      if (e < dc) { ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0; }   // trace.h:220-221, after the sweep
      trace_push<DYN, T>(p, ag, tr, e, dc);
    }
  }
  else
  {
    for (int j = 0; j < T; ++j) tbl[ag.pidx[j]] += dw;        // sequential: duplicates accumulate
    if (p.trace_kind != 0)
    {
      // Representation::update(trace, ...) (representation.h:79-83): newest -> oldest
      double weight = 1.;
      for (int k = 0; k < ag.len && weight > 0.001; ++k)
      {
        const int pos = ring_pos(ag.head, k, cap);
        const double inc = weight * ad * e;
        for (int j = 0; j < T; ++j)
        {
          uint32_t ix = tr[((size_t)pos * T + j) * n];
          if (ix != kInvalidIndex) { tbl[ix] += inc; ++swept; }
        }
        weight *= e;
      }
      // EnumeratedTrace::add (trace.h:215-234 replacing, 244-261 accumulating)
      if (e < dc) { ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0; }
      if (p.trace_kind == 1)
        for (int k = 0; k < ag.len; ++k)
        {
          const int pos = ring_pos(ag.head, k, cap);
          for (int j = 0; j < T; ++j)
          {
            uint32_t ix = tr[((size_t)pos * T + j) * n];
            bool hit = false;
#pragma unroll
            for (int jj = 0; jj < T; ++jj) hit |= (ix == ag.pidx[jj]);
            if (hit) tr[((size_t)pos * T + j) * n] = kInvalidIndex;
          }
        }
      trace_push<DYN, T>(p, ag, tr, e, dc);
    }
  }
  ag.swept += (uint64_t)swept;
  o_qp = qp; o_target = target; o_delta = delta; o_swept = swept;
}

}  // namespace grlcu
