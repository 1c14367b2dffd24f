// grlcu_coop.cuh -- the fused loop with SIXTEEN lanes per agent (one per tiling) and L2-sized cohorts.
//
// Why: with one thread per agent (k_run) all agents of a context are resident at once, one step of all of them
// touches ~0.8 GB of distinct lines, nothing is ever found in L2 again and every table access is a DRAM request
// (the kernel sits on the measured request ceiling).  A third of those requests are repeats
// of what the same agent touched one step earlier (the 16 words of p, ~10 of the 48 words of Q(s',.)), and the
// dense trace-cache stream is the same ~4 KB per agent every step.  If only C agents are in flight and
// C * (64 lines * 128 B + 4 KB) fits in L2, those repeats and the whole dense stream become L2 hits.  One thread per
// agent cannot do that: C = 4096 agents is 28 threads per SM.  Sixteen lanes per agent can: lane j owns tiling j --
// its tile hash, its table words, column j of every trace entry -- so C agents are 16 C threads and one agent-step is
// ~10x shorter.  The agents of a context are processed cohort after cohort, n_steps each.
//
// Everything an agent owns that is not per tiling (state, RK4, draws, counters) is computed redundantly and
// identically by its 16 lanes.  Sums over tilings keep the reference's order (linear.cpp:147-151): the 16 words
// are passed around with shuffles and added in tiling order by every lane.
//
// The cache is the one of grlcu_cached.cuh (same slots, same sentinels, same arithmetic) in an agent-major layout:
// cv[la][pos*T + j], ring[la][pos*T + j], vmask[la][pos] for the la-th agent of the cohort, so a trace entry is one
// 128-byte line.  Cache<T> with n = 1 addresses it, which lets the rare cases -- duplicate indices inside a
// projection or inside a live trace entry -- run the tested single-thread code of grlcu_cached.cuh on lane 0.
#pragma once
#include "grlcu_agent.cuh"
#include "grlcu_cached.cuh"
#include "grlcu_kernels.cuh"

namespace grlcu {

constexpr int kLanes = 16;            // tilings of the instantiated kernels (T)

struct CoopArrays
{
  double   *cv;       // [cohort][cap*T]
  uint32_t *ring;     // [cohort][cap*T]
  uint32_t *vmask;    // [cohort][cap]
};

// Lane geometry: L lanes per agent, each owning TPL = T / L consecutive tilings (j = lane*TPL + t).
template <int L> struct Geo
{
  static constexpr int T = kLanes, TPL = T / L;
  static_assert(L == 2 || L == 4 || L == 8 || L == 16, "lanes per agent");
};

// the T words of a projection added in tiling order, starting from 0 like linear.cpp:147-151 (every lane gets the sum)
template <int L>
__device__ __forceinline__ double ordered_sum(uint32_t gm, const double (&w)[Geo<L>::TPL])
{
  double r = 0.;
#pragma unroll
  for (int l = 0; l < L; ++l)
  {
    double mine = r;
#pragma unroll
    for (int t = 0; t < Geo<L>::TPL; ++t) mine += w[t];
    r = __shfl_sync(gm, mine, l, L);
  }
  return r;
}

// true if any two of the agent's T indices are equal
template <int L>
__device__ __forceinline__ bool group_has_duplicates(uint32_t gm, int lane, const uint32_t (&x)[Geo<L>::TPL])
{
  constexpr int TPL = Geo<L>::TPL;
  bool dup = false;
#pragma unroll
  for (int t = 1; t < TPL; ++t)
#pragma unroll
    for (int u = 0; u < t; ++u) dup |= (x[t] == x[u]);
#pragma unroll
  for (int k = 1; k < L; ++k)
#pragma unroll
    for (int u = 0; u < TPL; ++u)
    {
      const uint32_t y = __shfl_sync(gm, x[u], (lane + k) & (L - 1), L);
#pragma unroll
      for (int t = 0; t < TPL; ++t) dup |= (x[t] == y);
    }
  return (__ballot_sync(gm, dup) & gm) != 0;
}

// Replicated per-agent state (identical in the L lanes) + the lane's own tilings of the previous projection.
template <int DYN, int L>
struct CoopAgent
{
  static constexpr int S = Dyn<DYN>::S;
  double   x[S];
  double   action;
  uint32_t pidx[Geo<L>::TPL];       // lane l: indices of tilings l*TPL .. l*TPL+TPL-1 of p
  int32_t  len, head, flags;
  uint64_t dupmask;
  double   total_decay, total_reward;
  uint64_t act_ctr;
  double   alpha, e, epsilon;
  uint32_t dsteps, dswept, dtblw;
};

// Rare path: duplicate indices in p or in a live trace entry.  Lane 0 runs grlcu_cached.cuh's update on the
// agent-major cache (Cache<T> with n = 1); the other lanes wait and take the new scalars from it.
template <int DYN, int L>
__device__ __noinline__ void coop_update_serial(const Params &p, CoopAgent<DYN, L> &ag, const Cache<kLanes> c1, double *tbl,
                                                uint32_t gm, int lane, const double (&rawp)[Geo<L>::TPL], double reward,
                                                bool bootstrap, double qn)
{
  constexpr int TPL = Geo<L>::TPL;
  Agent<DYN, kLanes> f;
  double rp[kLanes];
#pragma unroll
  for (int j = 0; j < kLanes; ++j)
  {
    f.pidx[j] = __shfl_sync(gm, ag.pidx[j % TPL], j / TPL, L);
    rp[j] = __shfl_sync(gm, rawp[j % TPL], j / TPL, L);
  }
  f.flags = ag.flags; f.alpha = ag.alpha; f.e = ag.e; f.head = ag.head; f.len = ag.len; f.dupmask = ag.dupmask;
  f.total_decay = ag.total_decay; f.dswept = 0; f.dtblw = 0;
  __syncwarp(gm);
  if (lane == 0)
  {
    double qp, target, delta; int swept;
    sarsa_update_cached<DYN, kLanes>(p, f, c1, tbl, reward, bootstrap, qn, rp, qp, target, delta, swept);
  }
  __syncwarp(gm);
  ag.head = __shfl_sync(gm, f.head, 0, L);
  ag.len = __shfl_sync(gm, f.len, 0, L);
  ag.dupmask = __shfl_sync(gm, (unsigned long long)f.dupmask, 0, L);
  ag.total_decay = __shfl_sync(gm, f.total_decay, 0, L);
  ag.dswept += __shfl_sync(gm, f.dswept, 0, L);
  ag.dtblw += __shfl_sync(gm, f.dtblw, 0, L);
}

// Write the lane's columns of one entry back into the table.
template <int L>
__device__ __forceinline__ uint32_t coop_flush_entry(const Cache<kLanes> &c1, double *tbl, int pos, int lane)
{
  constexpr int TPL = Geo<L>::TPL;
  const uint32_t m = c1.mask(pos);
  uint32_t ix[TPL];
  double v[TPL];
#pragma unroll
  for (int t = 0; t < TPL; ++t) { ix[t] = c1.idx(pos, lane * TPL + t); v[t] = c1.val(pos * kLanes + lane * TPL + t); }
#pragma unroll
  for (int t = 0; t < TPL; ++t)
    if ((m >> (lane * TPL + t)) & 1u) tbl[ix[t]] = v[t];
  return (uint32_t)__popc(m);
}

// trace.clear() (sarsa.cpp:126-132 via TDAgent::start)
template <int DYN, int L>
__device__ __forceinline__ void coop_clear(const Params &p, CoopAgent<DYN, L> &ag, const Cache<kLanes> &c1, double *tbl, int lane)
{
  for (int k = 0; k < ag.len; ++k) ag.dtblw += coop_flush_entry<L>(c1, tbl, ring_pos(ag.head, k, p.trace_cap), lane);
  ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0;
}

// SARSAPredictor::criticize (sarsa.cpp:98-124), replacing trace, through the value cache: the arithmetic of
// sarsa_update_cached (grlcu_cached.cuh), TPL tilings per lane.  rawp: table[pidx] as read before any store of this step.
template <int DYN, int L>
__device__ __forceinline__ void coop_update(const Params &p, CoopAgent<DYN, L> &ag, const Cache<kLanes> &c1, double *tbl,
                                            uint32_t gm, int lane, const double (&rawp)[Geo<L>::TPL], double reward,
                                            bool bootstrap, double qn)
{
  constexpr int TPL = Geo<L>::TPL;
  if ((ag.flags & 8) || ag.dupmask != 0)
  {
    coop_update_serial<DYN, L>(p, ag, c1, tbl, gm, lane, rawp, reward, bootstrap, qn);
    return;
  }
  const int cap = p.trace_cap;
  double target = reward;
  if (bootstrap) target += p.gamma * qn;        // pow(gamma, 1)
  int sp[TPL];
  double np[TPL];
#pragma unroll
  for (int t = 0; t < TPL; ++t)
  {
    sp[t] = is_sentinel(rawp[t]) ? sentinel_slot(rawp[t]) : -1;
    np[t] = sp[t] >= 0 ? c1.val(sp[t]) : rawp[t];
  }
  const double qp = ordered_sum<L>(gm, np) / kLanes;
  const double delta = target - qp;
  const double dw = ag.alpha * (target - qp);   // linear.cpp:194
  const double e = ag.e;
  const double ad = ag.alpha * delta;
  const double dc = 0.01;
#pragma unroll
  for (int t = 0; t < TPL; ++t) np[t] += dw;
  // an index that an older entry holds: that slot receives p's write now and the sweep's increment below
#pragma unroll
  for (int t = 0; t < TPL; ++t)
    if (sp[t] >= 0) c1.val(sp[t]) = np[t];
  __syncwarp(gm);
  // update(trace, alpha*delta, gamma*lambda) (representation.h:79-83): newest -> oldest, each lane on its columns;
  // two entries per round so that their loads are in flight together
  {
    double weight = 1.;
    uint32_t swept = 0;
    int k = 0;
    while (k < ag.len && weight > 0.001)
    {
      const int pos0 = ring_pos(ag.head, k, cap);
      const double inc0 = weight * ad * e;
      weight *= e;
      const bool two = k + 1 < ag.len && weight > 0.001;
      const int pos1 = two ? ring_pos(ag.head, k + 1, cap) : pos0;
      const double inc1 = weight * ad * e;
      if (two) weight *= e;
      const uint32_t m0 = c1.mask(pos0), m1 = c1.mask(pos1);
      double v0[TPL], v1[TPL];
#pragma unroll
      for (int t = 0; t < TPL; ++t) { v0[t] = c1.val(pos0 * kLanes + lane * TPL + t); v1[t] = c1.val(pos1 * kLanes + lane * TPL + t); }
#pragma unroll
      for (int t = 0; t < TPL; ++t)
        if ((m0 >> (lane * TPL + t)) & 1u) c1.val(pos0 * kLanes + lane * TPL + t) = v0[t] + inc0;
      swept += (uint32_t)__popc(m0);
      if (two)
      {
#pragma unroll
        for (int t = 0; t < TPL; ++t)
          if ((m1 >> (lane * TPL + t)) & 1u) c1.val(pos1 * kLanes + lane * TPL + t) = v1[t] + inc1;
        swept += (uint32_t)__popc(m1);
      }
      k += two ? 2 : 1;
    }
    ag.dswept += swept;
  }
  __syncwarp(gm);
  // trace.add(p, e) (trace.h:215-234).  ssub: the old slot named by p's sentinel is invalidated, its weight moves
#pragma unroll
  for (int t = 0; t < TPL; ++t)
    if (sp[t] >= 0)
    {
      np[t] = c1.val(sp[t]);
      atomicAnd(&c1.mask(sp[t] / kLanes), ~(1u << (sp[t] % kLanes)));
    }
  __syncwarp(gm);
  if (e < dc) coop_clear<DYN, L>(p, ag, c1, tbl, lane);                // trace.h:220-221
  ag.head = ag.head + 1 == cap ? 0 : ag.head + 1;
#pragma unroll
  for (int t = 0; t < TPL; ++t)
  {
    const int slot = ag.head * kLanes + lane * TPL + t;
    c1.val(slot) = np[t];
    tbl[ag.pidx[t]] = make_sentinel(slot);
    c1.idx(ag.head, lane * TPL + t) = ag.pidx[t];
  }
  if (lane == 0) c1.mask(ag.head) = 0xFFFFu;
  ag.dtblw += kLanes;
  __syncwarp(gm);
  ag.len++;
  ag.dupmask <<= 1;
  ag.total_decay *= e;
  while (ag.total_decay < dc && ag.len > 1)
  {
    ag.total_decay /= e;
    ag.dtblw += coop_flush_entry<L>(c1, tbl, ring_pos(ag.head, ag.len - 1, cap), lane);
    ag.len--;
  }
}

template <int DYN, int A, int L, int MINB>
__global__ void __launch_bounds__(kMaxBlock, MINB)
k_run_coop(Params p, CoopArrays ca, int agent_begin, int agent_end, int64_t n_steps)
{
  constexpr int S = Dyn<DYN>::S, OBS = Dyn<DYN>::OBS, T = kLanes, TPL = Geo<L>::TPL;
  const int lane = threadIdx.x & (L - 1);
  const int la = (blockIdx.x * blockDim.x + threadIdx.x) / L;        // agent of the cohort
  const int a = agent_begin + la;
  if (a >= agent_end) return;                                         // the whole group leaves together
  const uint32_t gm = (L == 32 ? 0xFFFFFFFFu : ((1u << L) - 1u)) << ((threadIdx.x & 31 & ~(L - 1)));   // this group's lanes of the warp
  const uint64_t gid = (uint64_t)(p.first_id + a);
  double *tbl = p.table + (size_t)a * p.proj.memory;
  const size_t n = (size_t)p.n;
  const int cap = p.trace_cap;
  Cache<T> c1;
  c1.cv = ca.cv + (size_t)la * cap * T; c1.ring = ca.ring + (size_t)la * cap * T; c1.vmask = ca.vmask + (size_t)la * cap;
  c1.n = 1u;
  uint32_t kj[TPL], kact[A][TPL];
#pragma unroll
  for (int t = 0; t < TPL; ++t)
  {
    kj[t] = p.kj[lane * TPL + t];
#pragma unroll
    for (int aa = 0; aa < A; ++aa) kact[aa][t] = p.kact[aa][lane * TPL + t];
  }

  CoopAgent<DYN, L> ag;
#pragma unroll
  for (int i = 0; i < S; ++i) ag.x[i] = p.state[(size_t)i * n + a];
  ag.action = p.action[a];
#pragma unroll
  for (int t = 0; t < TPL; ++t) ag.pidx[t] = p.prev_idx[(size_t)(lane * TPL + t) * n + a];
  ag.len = p.trace_len[a]; ag.head = p.trace_head[a]; ag.flags = p.flags[a]; ag.dupmask = p.dupmask[a];
  ag.total_decay = p.total_decay[a]; ag.total_reward = p.total_reward[a];
  ag.act_ctr = p.act_ctr[a];
  ag.alpha = p.alpha_a ? p.alpha_a[a] : p.alpha;
  ag.e = p.gamma * (p.lambda_a ? p.lambda_a[a] : p.lambda);
  ag.epsilon = p.epsilon_a ? p.epsilon_a[a] : p.epsilon;
  ag.dsteps = 0; ag.dswept = 0; ag.dtblw = 0;
  uint64_t tt = p.tt[a], episode_ctr = p.episode_ctr[a], test_act_ctr = p.test_act_ctr[a];
  uint32_t episodes = 0;
  double last_return = 0, last_test_return = 0, return_sum = p.return_sum[a];
  bool have_return = false, have_test_return = false;

  // open: the plain trace ring of this agent into the agent-major cache (grlcu_cached.cuh, cache_open_core)
  for (int k = 0; k < ag.len; ++k)
  {
    const int pos = ring_pos(ag.head, k, cap);
#pragma unroll
    for (int t = 0; t < TPL; ++t) c1.idx(pos, lane * TPL + t) = p.trace_idx[((size_t)pos * T + lane * TPL + t) * n + a];
  }
  __syncwarp(gm);
  {
    uint32_t st = 0;
    if (lane == 0) st = cache_open_core<T>(c1, tbl, cap, ag.head, ag.len, ag.dupmask);
    __syncwarp(gm);
    ag.dtblw += __shfl_sync(gm, st, 0, L);
  }

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
      test = p.test_interval >= 0 && tt % (uint64_t)(p.test_interval + 1) == (uint64_t)p.test_interval;
      const uint64_t ep = episode_ctr++;
      if (p.start_override && (ag.flags & 4))
      {
#pragma unroll
        for (int i = 0; i < S; ++i) ag.x[i] = p.start_override[(size_t)i * n + a];
      }
      else
      {
        Philox4 w = philox_for(p.seed, gid, STREAM_START, ep);
        task_start<DYN>(u53(w.w[0], w.w[1]), p.randomization, test ? 1 : 0, ag.x);
      }
      int term0;
      observe<DYN>(ag.x, p.timeout, obs, term0, p.end_stop_penalty);
      if (!test) coop_clear<DYN, L>(p, ag, c1, tbl, lane);    // predictor finalize -> trace clear
      ag.total_reward = 0;
      ag.flags = 1 | (test ? 2 : 0);
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
    const bool act = terminal != 2, learn = !starting && !test;

    Draws d; d.u = 1.; d.r_explore = 0; d.r_tie = 0;
    if (act)
    {
      if (test) d = draws_from(philox_for(p.seed, gid, STREAM_TEST_ACT, test_act_ctr++));
      else d = draws_from(philox_for(p.seed, gid, STREAM_ACT, ag.act_ctr++));
    }

    uint32_t nidx[A][TPL];
    double rawp[TPL];
    double qn = 0;
    int ai = 0;
    if (act)
    {
      // TileCodingProjector::_project of (obs, a) for the lane's tilings (tile_coding.cpp:103-149), every action
      const uint32_t m = 0x5bd1e995u;
      int32_t qz[OBS];
#pragma unroll
      for (int i = 0; i < OBS; ++i) qz[i] = (int32_t)floor(obs[i] * p.proj.scaling[i]);
      double w[A][TPL];
#pragma unroll
      for (int t = 0; t < TPL; ++t)
      {
        const int j = lane * TPL + t;
        uint32_t h = 449u ^ (uint32_t)(OBS + 2);
#pragma unroll
        for (int i = 0; i < OBS; ++i)
        {
          int32_t cc = qz[i] - safe_mod(qz[i] - j * (1 + 2 * i), T);
          if (p.proj.wrap[i] != 0) cc = safe_mod(cc, p.proj.wrap[i]);
          h = (h * m) ^ murmur_mix((uint32_t)cc);
        }
#pragma unroll
        for (int aa = 0; aa < A; ++aa)
        {
          uint32_t g = (h * m) ^ kact[aa][t];
          g = (g * m) ^ kj[t];
          g ^= g >> 13; g *= m; g ^= g >> 15;
          nidx[aa][t] = reduce_index(p, g);
        }
      }
#pragma unroll
      for (int aa = 0; aa < A; ++aa)
#pragma unroll
        for (int t = 0; t < TPL; ++t) w[aa][t] = ld_tbl(tbl + nidx[aa][t]);
      if (learn)
      {
#pragma unroll
        for (int t = 0; t < TPL; ++t) rawp[t] = ld_tbl(tbl + ag.pidx[t]);
      }
      // LinearRepresentation::read (linear.cpp:136-184) through the directory
      double q[A];
#pragma unroll
      for (int aa = 0; aa < A; ++aa)
      {
#pragma unroll
        for (int t = 0; t < TPL; ++t)
          if (is_sentinel(w[aa][t])) w[aa][t] = c1.val(sentinel_slot(w[aa][t]));
      }
#pragma unroll
      for (int aa = 0; aa < A; ++aa) q[aa] = ordered_sum<L>(gm, w[aa]) / T;
      ai = sample_action<A>(q, d, ag.epsilon, test, p.softmax_tau);
      qn = bootstrap_value<A>(q, ai, p.target_kind, p.target_epsilon);
    }
    else if (learn)
    {
#pragma unroll
      for (int t = 0; t < TPL; ++t) rawp[t] = ld_tbl(tbl + ag.pidx[t]);
    }

    if (learn) coop_update<DYN, L>(p, ag, c1, tbl, gm, lane, rawp, reward, act, qn);

    if (act)
    {
#pragma unroll
      for (int t = 0; t < TPL; ++t)
      {
        uint32_t x = nidx[0][t];
#pragma unroll
        for (int aa = 1; aa < A; ++aa) if (aa == ai) x = nidx[aa][t];
        ag.pidx[t] = x;
      }
      ag.flags = (ag.flags & ~8) | (group_has_duplicates<L>(gm, lane, ag.pidx) ? 8 : 0);
      ag.action = p.actions[ai];
    }
    if (!starting)
    {
      if (!test) ag.dsteps++;
      if (terminal)
      {
        // online_learning.cpp:223-262 bookkeeping at the end of an episode
        ag.flags = 0;
        tt++; episodes++;
        if (test) { last_test_return = ag.total_reward; have_test_return = true; }
        else { last_return = ag.total_reward; have_return = true; return_sum += ag.total_reward; }
      }
      ++step;
    }
  }

  // close: back to the plain representation (cache_close_core), then the ring in its SoA home
  __syncwarp(gm);
  {
    uint32_t st = 0;
    if (lane == 0) st = cache_close_core<T>(c1, tbl, cap, ag.head, ag.len);
    __syncwarp(gm);
    ag.dtblw += __shfl_sync(gm, st, 0, L);
  }
  for (int k = 0; k < ag.len; ++k)
  {
    const int pos = ring_pos(ag.head, k, cap);
#pragma unroll
    for (int t = 0; t < TPL; ++t) p.trace_idx[((size_t)pos * T + lane * TPL + t) * n + a] = c1.idx(pos, lane * TPL + t);
  }
#pragma unroll
  for (int t = 0; t < TPL; ++t) p.prev_idx[(size_t)(lane * TPL + t) * n + a] = ag.pidx[t];
  if (lane == 0)
  {
#pragma unroll
    for (int i = 0; i < S; ++i) p.state[(size_t)i * n + a] = ag.x[i];
    p.action[a] = ag.action;
    p.trace_len[a] = ag.len; p.trace_head[a] = ag.head; p.flags[a] = ag.flags; p.dupmask[a] = ag.dupmask;
    p.total_decay[a] = ag.total_decay; p.total_reward[a] = ag.total_reward;
    p.act_ctr[a] = ag.act_ctr;
    p.ss[a] += ag.dsteps; p.swept[a] += ag.dswept; p.tblw[a] += ag.dtblw;
    p.tt[a] = tt; p.episode_ctr[a] = episode_ctr; p.test_act_ctr[a] = test_act_ctr;
    p.episodes[a] += episodes;
    if (have_test_return) p.last_test_return[a] = last_test_return;
    if (have_return) p.last_return[a] = last_return;
    p.return_sum[a] = return_sum;
  }
}

}  // namespace grlcu
