// grlcu_agent.cuh -- one agent-step (online_learning.cpp:186-221) for one agent per thread.
//
// Layout in HBM (all per-context, structure-of-arrays over agents so that a warp's 32
// agents read/write consecutive addresses for everything except the hashed tables):
//   table      double  [n][memory]          per-agent hashed weight table (agent-major)
//   state      double  [S][n]               environment state
//   action     double  [n]                  action to apply at the next step
//   prev_idx   uint32  [T][n]               projection of (prev_obs, prev_action)
//   trace_idx  uint32  [n/32][cap*T][32]    enumerated trace ring (0xFFFFFFFF = invalidated), blocked by warp (warp_rows)
//   trace_len/head int [n], total_decay double [n]
//   counters / returns [n]
// Citations: /root/reference/base.
#pragma once
#include <math_constants.h>
#include "grlcu_device.cuh"

namespace grlcu {

// Physical table layout ("tile-major").  The reference addresses its table with the raw hash,
// x = h % memory (tile_coding.h:116-118), so the A words of Q(s', .) that belong to one tile of one tiling are A
// unrelated DRAM requests.  x stays the LOGICAL address (what replay reports, what grlcu_get_table / set_table and
// checkpoints use); word x is STORED at physical slot perm[x], one permutation per context shared by all its agents
// (same hash, same memory => the same collisions for every agent), built on the host by grlcu_create:
//   * the tiles inside the task's observation box are enumerated cell by cell (cell = the T-aligned coarse tile
//     coordinates k_i of every observation dimension), tiling by tiling, action by action; the A action words of one
//     (cell, tiling) group get consecutive slots, padded to one 32-byte sector: slot = ((cell * T) + j) * 4 + a;
//   * a word whose index x was already placed by an earlier tile (a hash collision) simply shares that earlier slot
//     (slotmap[] names it; slotmap[] of a word that owns its slot is that slot): semantics unchanged by construction;
//   * indices no tile of the box maps to fill the holes and the tail.
// The fused kernel computes the group from the tile coordinates (no hash) and takes the slots of its A words from
// slotmap[] with asynchronous global -> shared copies (stash_async); the MurmurHash2 path + perm[] serve observations
// outside the box.
struct Geo
{
  int32_t  enabled;             // 1: slots from tile coordinates inside the box (perm[] outside); 2: perm[] only (visit-ordered layout)
  int32_t  kn[kMaxDims];        // coarse cells per observation dimension
  int32_t  koff[kMaxDims];      // k + koff in [0, kn) inside the box (unwrapped dimensions)
  int32_t  wrapped[kMaxDims];   // dimension wraps (wrap = kn * T): k mod kn
  const uint32_t *perm;         // [memory] logical index -> physical slot
  const uint32_t *iperm;        // [tstride] physical slot -> logical index (kInvalidIndex: padding)
  const uint32_t *slotmap;      // [ncells * T * 4] physical slot of every in-box tile word
  const uint32_t *gflags;       // 4 bits per (cell, tiling) group: bit a = action a's word is in the group's own sector (layout statistics; the kernels read slotmap[] only)
  int32_t  *mismatch;           // replay kernels: geometric slot != perm[hash index] seen (must stay 0)
  // pilot run of grlcu_create (visit-ordered layout): every projected group is logged as (hash state after the
  // observation coordinates) << 8 | tiling
  uint64_t *pilot_keys;         // [n][pilot_cap]
  uint32_t *pilot_count;        // [n]
  uint32_t  pilot_cap;
};

struct Params
{
  int32_t  n;                 // agents in this context
  int64_t  first_id;          // global id of agent 0 (Philox stream)
  uint64_t seed;
  double   tau; int32_t isteps; double timeout, randomization;
  int32_t  end_stop_penalty;  // task/cart_pole/swingup end_stop_penalty (cart_pole.cpp:177,196)
  Projector proj;
  Geo      geo;
  uint32_t tstride;           // doubles per agent in `table` (memory, or more with the padded tile-major layout)
  uint32_t kact[8][16];      // pre-mixed MurmurHash2 word of the action coordinate, per (action, tiling)
  uint32_t kj[16];           // pre-mixed word of the tiling number
  uint32_t mem_mask;         // memory-1 if memory is a power of two, else 0
  uint64_t mem_magic;        // 2^64/memory + 1 (exact 32-bit remainder by multiplication)
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
  int32_t  *flags;            // bit0 in_episode, bit1 test, bit2 start override pending, bit3 prev projection has duplicate indices,
                              // bit4 the agent's trace value cache is live (its table holds directory sentinels, grlcu_cached.cuh)
  uint64_t *dupmask;          // bit k: trace entry of age k holds duplicate indices (fast paths: cap <= 64)
  int32_t  force_slow;        // 1: always take the sequential update path (A/B testing)
  int32_t  mode;              // 0: trace value cache where eligible, 1: uncached
  int32_t  target_kind;       // bootstrap value: 0 Q(s',a') (SARSA), 1 max_a Q(s',a) (Q-learning), 2 policy value (expected SARSA)
  double   target_epsilon;    // expected SARSA: target policy's epsilon (< 0: greedy sampler)
  double   softmax_tau;       // > 0: the learning policy's sampler is sampler/softmax with this temperature
  double   *cache_val;        // [n / 32][cap*T][32] cached trace weights, blocked by warp (grlcu_cached.cuh)
  uint32_t cblock;            // agents per launch cohort (grlcu_run)
  uint32_t *cache_mask;       // [n/32][cap][32] valid bits per trace entry, blocked by warp
  uint64_t *act_ctr, *test_act_ctr, *episode_ctr, *tt, *ss, *swept;
  uint64_t *tblw;             // random 8-byte table stores issued (roofline accounting)
  uint32_t *episodes;
  double   *total_reward, *last_return, *last_test_return, *return_sum;
  const double *alpha_a, *lambda_a, *epsilon_a;   // optional per-agent hyper-parameters
  const double *start_override;                    // optional [S][n] start states (run_host)
  // Per-agent lambda: launch order of the 32-agent chunks, sorted by lambda (chunk_map[k] = chunk that thread-chunk k of a
  // launch runs), so that a cohort launch holds agents with traces of similar length; NULL = identity
  const uint32_t *chunk_map;
  int32_t  *oob;              // replay kernels: a physical slot beyond the agent's table row was seen (must stay 0)
  // projector/tile_coding safe > 0 (tile_coding.h:116-148): per-agent collision-claim table and what goes with it
  int32_t  safe;              // 0 off, 1 claim on the predictor's projections, 2 claim on every projection
  int32_t  *claims;           // [n][memory] TileCodingProjector::indices_ (-1 = free)
  uint32_t *prev_hash;        // [T][n] tile hashes of (prev_obs, prev_action): the predictor projects them again, claiming
  uint32_t *pressure;         // [n] largest number of probe steps a lookup took (pressure_)
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

// What the predictor bootstraps from at s' (the policy's Q(s', .) re-used: no table write lies in between):
//   SARSAPredictor (sarsa.cpp:103-104): Q(s', a');
//   QPredictor (advantage.cpp:84-90): v = -inf; v = fmax(v, Q(s', a_k)) over the discretizer's options, in order;
//   ExpectedSARSAPredictor (sarsa.cpp:172-173) -> QPolicy::value (q.cpp:56-68): sum_k q_k * dist_k, dist from
//   GreedySampler::distribution (greedy.cpp:88-98: 1/man on the maxima) or EpsilonGreedySampler::distribution
//   (greedy.cpp:220-238, one epsilon, decay_ = 1: a 1 becomes 1 - eps, then every entry += eps / A).
template <int A>
__device__ __forceinline__ double bootstrap_value(const double (&q)[A], int ai, int kind, double target_epsilon)
{
  if (kind == 1)
  {
    double v = -CUDART_INF;
#pragma unroll
    for (int kk = 0; kk < A; ++kk) v = fmax(v, q[kk]);
    return v;
  }
  if (kind == 2)
  {
    int mai = 0, man = 1;
#pragma unroll
    for (int ii = 1; ii < A; ++ii)
    {
      if (q[ii] > q[mai]) { mai = ii; man = 1; }
      else if (q[ii] == q[mai]) man++;
    }
    double qm = q[0];
#pragma unroll
    for (int ii = 1; ii < A; ++ii) if (ii == mai) qm = q[ii];
    double v = 0;
#pragma unroll
    for (int ii = 0; ii < A; ++ii)
    {
      double d = (q[ii] == qm) ? 1. / man : 0.;
      if (target_epsilon >= 0)
      {
        if (d == 1) d = 1 - 1. * target_epsilon;
        d += 1. * target_epsilon / A;
      }
      v += q[ii] * d;
    }
    return v;
  }
  double qn = q[0];
#pragma unroll
  for (int aa = 1; aa < A; ++aa) if (aa == ai) qn = q[aa];
  return qn;
}

// SoftmaxSampler::distribution (softmax.cpp:61-101) + grl::sample(dist, 1.) (utils.h:227-240) on the uniform draw u.
// Out of line: only contexts configured with sampler/softmax ever call it (code size, see q_values).
template <int A>
__device__ __noinline__ int softmax_pick(const double *q, double tau, double u)
{
  double v[A], dist[A];
  double max_power = -1.7976931348623157e308;
#pragma unroll
  for (int ii = 0; ii < A; ++ii)
  {
    v[ii] = 0;
    const double pw = q[ii] / tau;
    max_power = (max_power < pw) ? pw : max_power;
  }
  const double min_power = max_power + -100.;
  const double center = (max_power + min_power) / 2.0;
  double sum = 0;
#pragma unroll
  for (int ii = 0; ii < A; ++ii)
  {
    double pw = q[ii] / tau;
    if (pw > min_power) { pw -= center; v[ii] = exp(pw); sum += v[ii]; dist[ii] = 1; }
    else dist[ii] = 0;
  }
#pragma unroll
  for (int ii = 0; ii < A; ++ii) dist[ii] *= v[ii] / sum;
  const double r = u * 1.;
  sum = 0;
#pragma unroll
  for (int ii = 0; ii < A; ++ii)
  {
    sum += dist[ii];
    if (r < sum) return ii;
  }
  return A - 1;
}

// greedy.cpp:47-61 findmax + 63-86 / 151-218 sample (epsilon_.size() == 1, decay_ = 1); softmax_tau > 0: the learning
// policy samples with sampler/softmax instead (test episodes stay greedy)
template <int A>
__device__ __forceinline__ int sample_action(const double (&q)[A], const Draws &d, double epsilon, bool test,
                                             double softmax_tau = 0.)
{
  if (!test && softmax_tau > 0.) return softmax_pick<A>(q, softmax_tau, d.u);
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

// 4-byte global -> shared copy that occupies no register and does not stall the issuing thread (LDGSTS); a thread
// sees its own copies after stash_wait().  Used for table-layout lookups whose only consumer is the slot stash.
__device__ __forceinline__ void stash_async(uint32_t *smem_dst, const uint32_t *src)
{
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void stash_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// TileCodingProjector::getFeatureLocation with a claim table (tile_coding.h:116-148): linear probing from h % memory
// until the entry is h or free; `indices_[ii] != h` compares the int32 entry with the unsigned hash, `!= -1` as int.
__device__ __noinline__ uint32_t probe_location(int32_t *claims, uint32_t memory, uint32_t h, bool claim, uint32_t *pressure)
{
  uint32_t ii = h % memory, collisions = 0;
  // (a full table makes the reference probe forever; here the walk stops after one lap so that a mis-sized table cannot
  // hang the GPU -- the pressure counter then reads `memory`)
  while ((uint32_t)claims[ii] != h && claims[ii] != -1 && collisions < memory)
  {
    ++collisions;
    if (++ii >= memory) ii = 0;
  }
  if (collisions > *pressure) *pressure = collisions;
  if (claim) claims[ii] = (int32_t)h;
  return ii;
}

constexpr int kMaxBlock = 64;     // threads per block of the fused kernels (shared index stash is sized for it)

// h % memory (tile_coding.h:118) without a division: mask for powers of two, else the exact
// multiply-high remainder of Lemire, Kaser & Kurz (2019): r = hi64((M*h mod 2^64) * d), M = 2^64/d + 1.
__device__ __forceinline__ uint32_t reduce_index(const Params &p, uint32_t h)
{
  if (p.mem_mask) return h & p.mem_mask;
  const uint64_t low = p.mem_magic * (uint64_t)h;
  return (uint32_t)__umul64hi(low, (uint64_t)(uint32_t)p.proj.memory);
}

__device__ __forceinline__ uint32_t murmur_mix(uint32_t k)
{
  const uint32_t m = 0x5bd1e995u;
  k *= m; k ^= k >> 24; k *= m;
  return k;
}

// Per-thread view of one agent's persistent state.  Only what every step touches lives in
// registers; episode-level counters and returns are read/updated in global memory when an
// episode starts or ends (once per ~100 steps).
template <int DYN, int T>
struct Agent
{
  static constexpr int S = Dyn<DYN>::S;
  double   x[S];
  double   action;
  uint32_t pidx[T];
  int32_t  len, head, flags;
  uint64_t dupmask;
  double   total_decay, total_reward;
  uint64_t act_ctr;
  double   alpha, e, epsilon;       // e = gamma*lambda = pow(gamma*lambda, tau) for tau = 1
  uint32_t dsteps, dswept;          // learning steps / swept indices since load()
  uint32_t dtblw;                   // table stores since load()

  __device__ __forceinline__ void load(const Params &p, int a)
  {
#pragma unroll
    for (int i = 0; i < S; ++i) x[i] = p.state[(size_t)i * p.n + a];
    action = p.action[a];
#pragma unroll
    for (int j = 0; j < T; ++j) pidx[j] = p.prev_idx[(size_t)j * p.n + a];
    len = p.trace_len[a]; head = p.trace_head[a]; flags = p.flags[a]; dupmask = p.dupmask[a];
    total_decay = p.total_decay[a]; total_reward = p.total_reward[a];
    act_ctr = p.act_ctr[a];
    alpha = p.alpha_a ? p.alpha_a[a] : p.alpha;
    e = p.gamma * (p.lambda_a ? p.lambda_a[a] : p.lambda);
    epsilon = p.epsilon_a ? p.epsilon_a[a] : p.epsilon;
    dsteps = 0; dswept = 0; dtblw = 0;
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
    p.act_ctr[a] = act_ctr;
    p.ss[a] += dsteps; p.swept[a] += dswept; p.tblw[a] += dtblw;
  }
  // online_learning.cpp:223-262 bookkeeping at the end of an episode
  __device__ __forceinline__ void episode_end(const Params &p, int a, bool test)
  {
    flags &= (16 | 4);       // keep: cache live, start-state override pending
    p.tt[a] += 1;
    p.episodes[a] += 1;
    if (test) p.last_test_return[a] = total_reward;
    else { p.last_return[a] = total_reward; p.return_sum[a] += total_reward; }
  }
};

// QPolicy::values (q.cpp:94-107): TileCodingProjector::_project of (obs, a) for every action
// (tile_coding.cpp:103-149) + LinearRepresentation::read (linear.cpp:136-184: sequential sum in tiling
// order, then / T).  The MurmurHash2 state after the observation coordinates is shared by the A actions
// (the action coordinate and the tiling number are the last two words and are pre-mixed on the host).
// PHYSICAL slots go to the shared-memory stash `sid` (element (aa*T+j) at sid[(aa*T+j)*kMaxBlock]); with GEO they
// come from the tile coordinates (see struct Geo), otherwise slot = index.  hash_out / index_out (replay): the
// reference's hash and logical index h % memory of every word.
// The gathers are issued in two groups (actions 0..A-2, then action A-1 together with the T entries of
// the previous projection p) to bound the registers holding loads in flight.
// SAFE (projector safe > 0): the location of a hash comes from the agent's claim table (probe_location; the policy's
// multi-projection claims only with safe > 1, tile_coding.h:66-72) and the hashes are kept in the stash rows
// [A*T, 2*A*T) for the predictor, which projects again, claiming (agent_core).
template <int DYN, int T, int A, bool GEO, bool SAFE, typename Resolve>
__device__ __forceinline__ void q_values(const Params &p, const double *tbl, const double (&obs)[Dyn<DYN>::OBS],
                                         uint32_t *sid, double (&q)[A], const uint32_t (&pidx)[T],
                                         double (&vp)[T], bool with_p, uint32_t *hash_out, uint32_t *index_out,
                                         int n, int a, Resolve resolve)
{
  constexpr int OBS = Dyn<DYN>::OBS;
  const uint32_t m = 0x5bd1e995u;
  // Phase 1: the A*T slots into the stash.  Pure integer work, rolled over the tilings on purpose: unrolled it
  // is ~100 KB of SASS per call site and the fused kernels are instruction-fetch bound.
  {
    int32_t qz[OBS];
#pragma unroll
    for (int i = 0; i < OBS; ++i) qz[i] = (int32_t)floor(obs[i] * p.proj.scaling[i]);
#pragma unroll 1
    for (int j = 0; j < T; ++j)
    {
      bool inbox = false;
      uint32_t group = 0;
      if (GEO && p.geo.enabled == 1)
      {
        // coarse tile coordinate of tiling j in dimension i: c_i = b + T*k with b = j(1+2i) mod T (tile_coding.cpp:122-131)
        inbox = true;
        uint32_t cell = 0;
#pragma unroll
        for (int i = 0; i < OBS; ++i)
        {
          int32_t k = (qz[i] - ((j * (1 + 2 * i)) & (T - 1))) >> (T == 16 ? 4 : T == 8 ? 3 : 5);
          k += p.geo.koff[i];
          if (p.geo.wrapped[i]) k = k < 0 ? k + p.geo.kn[i] : k >= p.geo.kn[i] ? k - p.geo.kn[i] : k;   // one period either way
          inbox &= (uint32_t)k < (uint32_t)p.geo.kn[i];            // anything else takes the hash path below
          cell = cell * (uint32_t)p.geo.kn[i] + (uint32_t)k;
        }
        group = cell * T + j;
      }
      if (GEO && inbox)
      {
        // slotmap[home] is home itself for a word that owns its slot and the earlier tile's slot for one that collided
        // (grlcu_layout.h).  The lookups go global -> shared asynchronously (no register, no wait): all A*T of a step are
        // in flight together and awaited once after the loop, where the rolled loop used to wait for a flag word and
        // then for the collided words' entries tiling by tiling (19 % of the stall samples, profiles/r02_ncu_pendulum.txt)
#pragma unroll
        for (int aa = 0; aa < A; ++aa) stash_async(&sid[(aa * T + j) * kMaxBlock], p.geo.slotmap + group * 4u + aa);
      }
      if (!GEO || !inbox || hash_out || index_out)
      {
        uint32_t h = 449u ^ (uint32_t)(OBS + 2);
#pragma unroll
        for (int i = 0; i < OBS; ++i)
        {
          int32_t c = qz[i] - safe_mod(qz[i] - j * (1 + 2 * i), T);
          if (p.proj.wrap[i] != 0) c = safe_mod(c, p.proj.wrap[i]);
          h = (h * m) ^ murmur_mix((uint32_t)c);
        }
        if (!GEO && p.geo.pilot_keys)
        {
          // one stream per agent, in the order the agent projects them (pilot_cap keys per agent)
          const uint32_t at = p.geo.pilot_count[a]++;
          if (at < p.geo.pilot_cap) p.geo.pilot_keys[(size_t)a * p.geo.pilot_cap + at] = ((uint64_t)h << 8) | (uint32_t)j;
        }
#pragma unroll
        for (int aa = 0; aa < A; ++aa)
        {
          uint32_t g = (h * m) ^ p.kact[aa][j];
          g = (g * m) ^ p.kj[j];
          g ^= g >> 13; g *= m; g ^= g >> 15;
          if (hash_out) hash_out[((size_t)aa * T + j) * n + a] = g;
          if (SAFE) { sid[(A * T + aa * T + j) * kMaxBlock] = g; continue; }    // located below, in the reference's order
          const uint32_t x = reduce_index(p, g);
          if (index_out) index_out[((size_t)aa * T + j) * n + a] = x;
          if (!GEO) sid[(aa * T + j) * kMaxBlock] = x;
          else
          {
            if (!inbox) stash_async(&sid[(aa * T + j) * kMaxBlock], p.geo.perm + x);       // awaited after the loop
            else
            {
              stash_wait();                                                              // replay only
              if (__ldg(p.geo.perm + x) != sid[(aa * T + j) * kMaxBlock]) *p.geo.mismatch = 1;   // must never happen
            }
          }
        }
      }
    }
  }
  if (GEO) stash_wait();
  if (SAFE)
  {
    // getFeatureLocation in the reference's order -- action by action, tiling by tiling (q.cpp:94-107 projects the
    // variants one after the other): with safe = 2 every lookup claims, so the order decides who gets which slot
    int32_t *claims = p.claims + (size_t)a * p.proj.memory;
#pragma unroll 1
    for (int k = 0; k < A * T; ++k)
    {
      const uint32_t x = probe_location(claims, (uint32_t)p.proj.memory, sid[(A * T + k) * kMaxBlock], p.safe > 1, p.pressure + a);
      sid[k * kMaxBlock] = x;
      if (index_out) index_out[(size_t)k * n + a] = x;
    }
  }
  // Phase 2: the gathers, issued in two groups (actions 0..A-2, then action A-1 together with p's T words) to
  // bound the registers holding loads in flight.  With GEO the A loads of one tiling normally hit ONE 32-byte
  // sector (L1 merges them into one request).
  {
    double w[A - 1][T];
#pragma unroll
    for (int aa = 0; aa < A - 1; ++aa)
#pragma unroll
      for (int j = 0; j < T; ++j) w[aa][j] = ld_tbl(tbl + sid[(aa * T + j) * kMaxBlock]);
#pragma unroll
    for (int aa = 0; aa < A - 1; ++aa)
    {
      double r = 0;
#pragma unroll
      for (int j = 0; j < T; ++j) r += resolve(w[aa][j]);
      q[aa] = r / T;    // clamp to +-DBL_MAX is the identity for finite values (linear.cpp:179-180)
    }
  }
  {
    double w[T];
#pragma unroll
    for (int j = 0; j < T; ++j) w[j] = ld_tbl(tbl + sid[((A - 1) * T + j) * kMaxBlock]);
    if (with_p)
    {
#pragma unroll
      for (int j = 0; j < T; ++j) vp[j] = ld_tbl(tbl + pidx[j]);
    }
    double r = 0;
#pragma unroll
    for (int j = 0; j < T; ++j) r += resolve(w[j]);
    q[A - 1] = r / T;
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

// ---- duplicate indices inside one projection (a hash collision between tilings) -------------------------
// Rare (about T*T/2 / memory per projection), so everything about them is kept out of line and rolled: the
// fused kernels are instruction-fetch bound when their code does not fit the instruction cache.

// bit j set: pidx[j] is the first occurrence of its index
__device__ __noinline__ uint32_t dup_owner_mask(const uint32_t *pidx, int T)
{
  uint32_t owner = (1u << T) - 1u;
#pragma unroll 1
  for (int j = 1; j < T; ++j)
#pragma unroll 1
    for (int jj = 0; jj < j; ++jj)
      if (((owner >> jj) & 1u) && pidx[jj] == pidx[j]) { owner &= ~(1u << j); break; }
  return owner;
}

// 4 bits per tiling jj: how many later entries duplicate owner jj (T <= 16)
__device__ __noinline__ uint64_t dup_counts(const uint32_t *pidx, uint32_t owner, int T)
{
  uint64_t counts = 0;
#pragma unroll 1
  for (int j = 1; j < T; ++j)
    if (!((owner >> j) & 1u))
#pragma unroll 1
      for (int jj = 0; jj < j; ++jj)
        if (((owner >> jj) & 1u) && pidx[jj] == pidx[j]) { counts += 1ull << (4 * jj); break; }
  return counts;
}

// 4 bits per tiling j: how many earlier tilings hold the same (valid) index
__device__ __noinline__ uint64_t dup_prior_counts(const uint32_t *ix, int T)
{
  uint64_t counts = 0;
#pragma unroll 1
  for (int j = 1; j < T; ++j)
    if (ix[j] != kInvalidIndex)
#pragma unroll 1
      for (int jj = 0; jj < j; ++jj)
        if (ix[jj] == ix[j]) counts += 1ull << (4 * j);
  return counts;
}

// the owner (first occurrence) of the index at tiling j
__device__ __noinline__ int dup_owner_of(const uint32_t *pidx, uint32_t owner, int j)
{
#pragma unroll 1
  for (int jj = 0; jj < j; ++jj)
    if (((owner >> jj) & 1u) && pidx[jj] == pidx[j]) return jj;
  return j;
}

// Install the projection chosen by the policy (read back from the index stash) as the agent's
// (prev_obs, prev_action) projection.
template <int DYN, int T>
__device__ __forceinline__ void set_prev_projection(Agent<DYN, T> &ag, const uint32_t *sid, int ai)
{
#pragma unroll
  for (int j = 0; j < T; ++j) ag.pidx[j] = sid[(ai * T + j) * kMaxBlock];
  ag.flags = (ag.flags & ~8) | (has_duplicates<T>(ag.pidx) ? 8 : 0);
}

// Per-agent rows of the trace arrays (trace_idx, cache_val, cache_mask) are blocked by warp: row r of agent a lives at
// base[(a / 32) * rows * 32 + r * 32 + a % 32].  A warp still reads one row as one contiguous segment, but the distance
// between an agent's rows is a compile-time constant (immediate offsets instead of a multiply-add per access) and the
// rows of one ring entry are contiguous in DRAM.
constexpr uint32_t kLanes = 32;
template <typename X>
__device__ __forceinline__ X *warp_rows(X *base, int a, size_t rows)
{
  return base + (size_t)((uint32_t)a >> 5) * rows * kLanes + ((uint32_t)a & 31u);
}
__host__ __device__ inline size_t warp_rows_agents(size_t n) { return (n + 31) / 32 * 32; }   // agents to allocate for

// EnumeratedTrace::add tail (trace.h:226-233 / 253-260): push p, total_decay *= decay, pop while
// total_decay < dc.  The ring only moves its head; popped entries are simply no longer counted.
template <int DYN, int T>
__device__ __forceinline__ void trace_push(const Params &p, Agent<DYN, T> &ag, uint32_t *tr, double e, double dc)
{
  const int cap = p.trace_cap;
  ag.head = ag.head + 1 == cap ? 0 : ag.head + 1;
  uint32_t *dst = tr + (size_t)ag.head * T * kLanes;
#pragma unroll
  for (int j = 0; j < T; ++j) dst[j * kLanes] = ag.pidx[j];
  ag.len++;
  ag.dupmask = (ag.dupmask << 1) | ((ag.flags & 8) ? 1ull : 0ull);
  ag.total_decay *= e;
  while (ag.total_decay < dc && ag.len > 1) { ag.total_decay /= e; ag.len--; }
}

// SARSAPredictor::criticize (sarsa.cpp:98-124) with tau = 1:
//   target = r [+ gamma*Q(s',a')]; delta = target - Q(p); write(p, target, alpha);
//   update(trace, alpha*delta, gamma*lambda); trace.add(p, gamma*lambda)
// vp[] holds the table values at p's indices, gathered by the caller before any store of this step.
//
// The reference applies every table increment as a sequential read-modify-write (p's T entries, then
// the trace entries newest to oldest).  Two code paths give bit-identical tables:
//  * slow path: exactly that sequence (needed whenever one address can receive two increments whose
//    order matters: duplicate indices inside a projection, or the accumulating trace);
//  * fast path (replacing trace, no duplicate indices in p or in any live entry): all valid trace
//    indices are then pairwise distinct (ssub removed every index an older entry shared with a newer
//    one, projection.h:93-104), so after p's T stores every remaining address gets exactly one
//    increment; the T gathers of an entry are independent and issued together, and ssub is done in the
//    same pass (same-tiling compare, then a 64-bit Bloom signature of p, then the exact T compares).
template <int DYN, int T>
__device__ __forceinline__ void sarsa_update(const Params &p, Agent<DYN, T> &ag, double *tbl, int a,
                                             double reward, bool bootstrap, double q_next, const double (&vp)[T],
                                             double &o_qp, double &o_target, double &o_delta, int &o_swept)
{
  const int cap = p.trace_cap;
  uint32_t *tr = warp_rows(p.trace_idx, a, (size_t)cap * T);   // element (pos, j) at tr[(pos*T + j)*kLanes]
  double target = reward;
  if (bootstrap) target += p.gamma * q_next;    // pow(gamma, 1)
  double qp = 0;
#pragma unroll
  for (int j = 0; j < T; ++j) qp += vp[j];
  qp /= T;
  const double delta = target - qp;
  // LinearRepresentation::write (linear.cpp:186-196): delta_w = alpha*(target - read(p))
  const double dw = ag.alpha * (target - qp);
  const double e = ag.e;
  const double ad = ag.alpha * delta;
  const double dc = p.trace_kind == 1 ? 0.01 : 0.0001;
  const bool fast = !p.force_slow && p.trace_kind != 2 && cap <= 64;
  int swept = 0;

  if (fast)
  {
    if (!(ag.flags & 8))
    {
#pragma unroll
      for (int j = 0; j < T; ++j) tbl[ag.pidx[j]] = vp[j] + dw;
    }
    else
    {
      // p holds duplicate indices (hash collision between tilings): the reference adds dw once per
      // occurrence, one after the other.  The last occurrence's store carries the full sum and lands last.
      uint32_t lp[T];
#pragma unroll
      for (int j = 0; j < T; ++j) lp[j] = ag.pidx[j];
#pragma unroll 1
      for (int j = 0; j < T; ++j) tbl[lp[j]] = ld_tbl(tbl + lp[j]) + dw;
    }
    if (p.trace_kind == 1)
    {
      uint64_t bloom = 0;
#pragma unroll
      for (int j = 0; j < T; ++j) bloom |= 1ull << (ag.pidx[j] & 63u);
      double weight = 1.;
      for (int k = 0; k < ag.len; ++k)
      {
        uint32_t *ent = tr + (size_t)ring_pos(ag.head, k, cap) * T * kLanes;
        const bool sw = weight > 0.001;           // representation.h:81
        const double inc = weight * ad * e;
        weight *= e;
        uint32_t ix[T];
        double v[T];
#pragma unroll
        for (int j = 0; j < T; ++j) ix[j] = ent[j * kLanes];
#pragma unroll
        for (int j = 0; j < T; ++j)
          if (sw && ix[j] != kInvalidIndex) v[j] = ld_tbl(tbl + ix[j]);
#pragma unroll
        for (int j = 0; j < T; ++j) v[j] += inc;
        if ((ag.dupmask >> k) & 1ull)
        {
          // duplicate indices inside this entry: one more sequential += inc per earlier occurrence
          uint32_t lx[T];
#pragma unroll
          for (int j = 0; j < T; ++j) lx[j] = ix[j];
          const uint64_t prior = dup_prior_counts(lx, T);
#pragma unroll
          for (int j = 1; j < T; ++j)
            for (int r = (int)((prior >> (4 * j)) & 15u); r > 0; --r) v[j] += inc;
        }
#pragma unroll
        for (int j = 0; j < T; ++j)
        {
          const uint32_t x = ix[j];
          if (x != kInvalidIndex)
          {
            if (sw) { tbl[x] = v[j]; ++swept; }
            bool hit = (x == ag.pidx[j]);                       // same tiling, same tile: the common case
            if (!hit && ((bloom >> (x & 63u)) & 1ull))
            {
#pragma unroll
              for (int jj = 0; jj < T; ++jj) hit |= (x == ag.pidx[jj]);
            }
            if (hit) ent[j * kLanes] = kInvalidIndex;                // ssub (projection.h:93-104)
          }
        }
      }
      if (e < dc) { ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0; }   // trace.h:220-221, after the sweep
      trace_push<DYN, T>(p, ag, tr, e, dc);
    }
  }
  else
  {
    for (int j = 0; j < T; ++j) tbl[ag.pidx[j]] = ld_tbl(tbl + ag.pidx[j]) + dw;   // sequential: duplicates accumulate
    if (p.trace_kind != 0)
    {
      // Representation::update(trace, ...) (representation.h:79-83): newest -> oldest
      double weight = 1.;
      for (int k = 0; k < ag.len && weight > 0.001; ++k)
      {
        const uint32_t *ent = tr + (size_t)ring_pos(ag.head, k, cap) * T * kLanes;
        const double inc = weight * ad * e;
        for (int j = 0; j < T; ++j)
        {
          uint32_t ix = ent[j * kLanes];
          if (ix != kInvalidIndex) { tbl[ix] = ld_tbl(tbl + ix) + inc; ++swept; }
        }
        weight *= e;
      }
      // EnumeratedTrace::add (trace.h:215-234 replacing, 244-261 accumulating)
      if (e < dc) { ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0; }
      if (p.trace_kind == 1)
        for (int k = 0; k < ag.len; ++k)
        {
          uint32_t *ent = tr + (size_t)ring_pos(ag.head, k, cap) * T * kLanes;
          for (int j = 0; j < T; ++j)
          {
            uint32_t ix = ent[j * kLanes];
            bool hit = false;
#pragma unroll
            for (int jj = 0; jj < T; ++jj) hit |= (ix == ag.pidx[jj]);
            if (hit) ent[j * kLanes] = kInvalidIndex;
          }
        }
      trace_push<DYN, T>(p, ag, tr, e, dc);
    }
  }
  ag.dswept += (uint32_t)swept;
  ag.dtblw += (uint32_t)(T + swept);
  o_qp = qp; o_target = target; o_delta = delta; o_swept = swept;
}

}  // namespace grlcu
