/* grl_oracle_agent.c -- oracle agent loop: Q policy, samplers, SARSA(lambda)/TD(lambda),
 * enumerated traces, actor-critic, experiment episode loop.
 * TEST INFRASTRUCTURE ONLY (see grl_oracle.h).  Citations: /root/reference/base. */
#define _DEFAULT_SOURCE
#define _XOPEN_SOURCE 600
#include "grl_oracle.h"
#include <math.h>
#include <float.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#define INVALID_INDEX 0xFFFFFFFFu   /* stands for (size_t)-1, projection.h:73-76 */

/* RNG stream ids for Philox mode (counter word 2) */
enum { STREAM_ACT = 0, STREAM_START = 1, STREAM_TABLE = 2, STREAM_TEST_ACT = 3,
       STREAM_ACTOR_TABLE = 4, STREAM_NOISE = 5 };

typedef struct
{
  int      len;                       /* entries, [0] oldest .. [len-1] newest */
  double   total_decay;
  double   decay[ORC_TRACE_STORE];
  uint32_t idx[ORC_TRACE_STORE][ORC_MAX_TILINGS];
} orc_trace;

struct orc_agent
{
  orc_config c;
  uint64_t id;
  int S;                              /* state dims */
  double alpha, lambda, epsilon;      /* per-agent (sweep) */
  double *table;                      /* Q (SARSA) or V (AC critic) */
  double *actor;                      /* AC actor table */
  int32_t *claims;                    /* safe > 0: TileCodingProjector::indices_ (tile_coding.cpp:51-56), -1 = free */
  uint64_t pressure;                  /* TileCodingProjector::pressure_ */
  /* glibc mode streams (utils.h:84-137) */
  struct drand48_data randgen;        /* thread-local RandGen: table init + task start */
  struct drand48_data sampler;        /* EpsilonGreedySampler::rand_ */
  struct drand48_data test_sampler;   /* GreedySampler::rand_ of test_agent */
  /* philox counters */
  uint64_t act_counter, test_act_counter, episode_counter;
  /* environment (modeled.cpp) */
  double state[ORC_MAX_STATE];
  double action;                      /* action to apply at the next env step */
  /* TDAgent state (agents/td.cpp) */
  double prev_obs[ORC_MAX_DIMS];
  double prev_action;
  uint32_t prev_idx[ORC_MAX_TILINGS]; /* projection of (prev_obs, prev_action); AC: critic projection of prev_obs */
  uint32_t prev_aidx[ORC_MAX_TILINGS]; /* AC: actor projection of prev_obs */
  double noise;                       /* AC: ActionPolicy::n_ */
  orc_trace trace;
  /* experiment (online_learning.cpp) */
  int in_episode, test;
  uint64_t tt, ss, episodes, swept_total;
  double total_reward, last_return, last_test_return;
};

/* ---- RNG ---------------------------------------------------------------- */

static double u53(uint32_t hi, uint32_t lo)
{
  uint64_t x = ((uint64_t)hi << 32) | lo;
  return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}

static void philox_for(const orc_agent *a, int stream, uint64_t ctr, uint32_t out[4])
{
  orc_philox((uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)stream, (uint32_t)(a->c.seed >> 32),
             (uint32_t)a->id, (uint32_t)a->c.seed ^ (uint32_t)(a->id >> 32), out);
}

void orc_glibc_seed(long seed) { srand48(seed); }

/* ---- LinearRepresentation (linear.cpp) ---------------------------------- */

/* linear.cpp:104-125 reset: params[ii] = rand->getUniform(init_min, init_max) */
static void table_init(orc_agent *a, double *tbl, int stream)
{
  const orc_config *c = &a->c;
  if (c->rng == ORC_RNG_GLIBC)
  {
    for (int32_t ii = 0; ii < c->memory; ++ii)
    {
      double r;
      drand48_r(&a->randgen, &r);
      tbl[ii] = c->init_min + r * (c->init_max - c->init_min);
    }
  }
  else
  {
    for (int32_t ii = 0; ii < c->memory; ii += 2)
    {
      uint32_t w[4];
      philox_for(a, stream, (uint64_t)(ii / 2), w);
      tbl[ii] = c->init_min + u53(w[0], w[1]) * (c->init_max - c->init_min);
      if (ii + 1 < c->memory)
        tbl[ii + 1] = c->init_min + u53(w[2], w[3]) * (c->init_max - c->init_min);
    }
  }
}

/* linear.cpp:136-184 read (IndexProjection, weights empty): sequential sum from 0,
 * divide by indices.size(), clamp to [omin, omax] */
static double table_read(const double *tbl, const uint32_t *idx, int T, double omin, double omax)
{
  double result = 0;
  for (int ii = 0; ii < T; ++ii)
    result += tbl[idx[ii]];
  result /= T;
  return fmin(fmax(result, omin), omax);
}

/* linear.cpp:198-216 update (limit_ = 1): skips invalidated indices */
static int table_update(double *tbl, const uint32_t *idx, int T, double delta, double omin, double omax)
{
  int n = 0;
  for (int ii = 0; ii != T; ++ii)
    if (idx[ii] != INVALID_INDEX)
    {
      tbl[idx[ii]] = fmin(fmax(tbl[idx[ii]] + delta, omin), omax);
      ++n;
    }
  return n;
}

/* linear.cpp:186-196 write: value = read(p); delta = alpha*(target-value); update(p, delta) */
static void table_write(double *tbl, const uint32_t *idx, int T, double target, double alpha,
                        double omin, double omax)
{
  double value = table_read(tbl, idx, T, omin, omax);
  double delta = alpha * (target - value);
  table_update(tbl, idx, T, delta, omin, omax);
}

/* ---- EnumeratedTrace (trace.h:135-263) ---------------------------------- */

static void trace_clear(orc_trace *t) { t->len = 0; t->total_decay = 1.; }

static void trace_pop_front(orc_trace *t)
{
  memmove(&t->decay[0], &t->decay[1], (size_t)(t->len - 1) * sizeof(double));
  memmove(&t->idx[0], &t->idx[1], (size_t)(t->len - 1) * sizeof(t->idx[0]));
  t->len--;
}

/* projection.h:93-104 ssub (weights empty) */
static void ssub(uint32_t *lhs, const uint32_t *rhs, int T)
{
  for (int ii = 0; ii < T; ++ii)
    for (int jj = 0; jj < T; ++jj)
      if (lhs[ii] == rhs[jj])
        lhs[ii] = INVALID_INDEX;
}

/* trace.h:215-234 (replacing) / 244-261 (accumulating) */
static void trace_add(orc_trace *t, int kind, const uint32_t *idx, int T, double decay)
{
  const double dc = kind == ORC_TRACE_REPLACING ? 0.01 : 0.0001;
  if (decay < dc)
    trace_clear(t);
  if (kind == ORC_TRACE_REPLACING)
    for (int ii = 0; ii < t->len; ++ii)
      ssub(t->idx[ii], idx, T);
  if (t->len == ORC_TRACE_STORE) abort();
  memcpy(t->idx[t->len], idx, (size_t)T * sizeof(uint32_t));
  t->decay[t->len] = decay;
  t->len++;
  t->total_decay *= decay;
  while (t->total_decay < dc && t->len > 1)
  {
    t->total_decay /= t->decay[0];
    trace_pop_front(t);
  }
}

/* representation.h:79-83: newest -> oldest while weight > 0.001;
 * update(projection, weight*delta*e); iterator weight_ *= decay of the entry just left
 * (trace.h:168-176) */
static void trace_sweep(const orc_trace *t, double *tbl, int T, double delta, double e,
                        double omin, double omax, int *swept, int *entries)
{
  double weight = 1.;
  int n = 0, m = 0;
  for (int k = 0; k < t->len && weight > 0.001; ++k)
  {
    int pos = t->len - k - 1;
    n += table_update(tbl, t->idx[pos], T, weight * delta * e, omin, omax);
    weight *= t->decay[pos];
    ++m;
  }
  if (swept) *swept = n;
  if (entries) *entries = m;
}

/* ---- samplers (greedy.cpp) ---------------------------------------------- */

typedef struct { double u; uint32_t r_explore, r_tie; int explored; } act_draws;

/* greedy.cpp:47-61 */
static void findmax(const double *values, int n, int *mai, int *man)
{
  *mai = 0; *man = 1;
  for (int ii = 1; ii < n; ++ii)
  {
    if (values[ii] > values[*mai]) { *mai = ii; *man = 1; }
    else if (values[ii] == values[*mai]) (*man)++;
  }
}

/* tie walk, greedy.cpp:77-85 / 206-214 */
static int tie_walk(const double *values, int mai, int r)
{
  int ii = mai;
  for (int jj = r; jj >= 0; ++ii)
    if (values[ii] == values[mai])
      --jj;
  return ii - 1;
}

/* raw 31-bit integer for getInteger(): lrand48() (utils.h:127-130) or a Philox word */
static uint32_t draw_int(const orc_agent *a, uint32_t philox_word)
{
  if (a->c.rng == ORC_RNG_GLIBC) return (uint32_t)lrand48();
  return philox_word >> 1;
}

/* SoftmaxSampler::distribution (softmax.cpp:61-101) followed by grl::sample(dist, 1.) (utils.h:227-240) on the
 * uniform draw u */
static int softmax_pick(const double *values, int n, double tau, double u)
{
  double v[ORC_MAX_ACTIONS], dist[ORC_MAX_ACTIONS];
  const double threshold = -100;
  double max_power = -DBL_MAX;
  for (int ii = 0; ii < n; ++ii)
  {
    v[ii] = 0;
    double p = values[ii] / tau;
    max_power = (max_power < p) ? p : max_power;
  }
  double min_power = max_power + threshold;
  double center = (max_power + min_power) / 2.0;
  double sum = 0;
  for (int ii = 0; ii < n; ++ii)
  {
    double p = values[ii] / tau;
    if (p > min_power)
    {
      p -= center;
      v[ii] = exp(p);
      sum += v[ii];
      dist[ii] = 1;
    }
    else
      dist[ii] = 0;
  }
  for (int ii = 0; ii < n; ++ii) dist[ii] *= v[ii] / sum;
  double r = u * 1.;
  sum = 0;
  for (int ii = 0; ii < n; ++ii)
  {
    sum += dist[ii];
    if (r < sum) return ii;
  }
  return n - 1;
}

/* EpsilonGreedySampler::sample (greedy.cpp:151-218, epsilon_.size()==1, decay_ = 1) or SoftmaxSampler::sample
 * (softmax.cpp:48-59) when test == 0, GreedySampler::sample (greedy.cpp:63-86) when test != 0. */
static int sample(orc_agent *a, const double *q, int test, act_draws *d)
{
  const int A = a->c.n_actions;
  uint32_t w[4] = {0, 0, 0, 0};
  int mai, man;
  memset(d, 0, sizeof(*d));
  findmax(q, A, &mai, &man);

  if (a->c.rng == ORC_RNG_PHILOX)
  {
    if (test) philox_for(a, STREAM_TEST_ACT, a->test_act_counter++, w);
    else      philox_for(a, STREAM_ACT, a->act_counter++, w);
  }

  if (!test && a->c.sampler == ORC_SAMPLER_SOFTMAX)
  {
    /* one uniform from RandGen (the thread's generator, not a sampler-owned Rand: utils.h:229) */
    double r;
    if (a->c.rng == ORC_RNG_GLIBC) drand48_r(&a->randgen, &r);
    else r = u53(w[0], w[1]);
    d->u = r;
    d->explored = 1;                           /* *at = atExploratory, always (softmax.cpp:54-55) */
    return softmax_pick(q, A, a->c.softmax_tau, r);
  }

  if (!test)
  {
    double r;
    if (a->c.rng == ORC_RNG_GLIBC) drand48_r(&a->sampler, &r);
    else r = u53(w[0], w[1]);
    d->u = r;
    if (r < 1. * a->epsilon)
    {
      d->r_explore = draw_int(a, w[2]);
      d->explored = 1;
      return (int)(d->r_explore % (uint32_t)A);
    }
  }
  if (man > 1)
  {
    d->r_tie = draw_int(a, w[3]);
    return tie_walk(q, mai, (int)(d->r_tie % (uint32_t)man));
  }
  return mai;
}

/* ---- TileCodingProjector::getFeatureLocation (tile_coding.h:116-148) ------ */
static uint32_t feature_location(orc_agent *a, uint32_t h, int claim)
{
  const uint32_t memory = (uint32_t)a->c.memory;
  uint32_t ii = h % memory;
  if (a->claims)
  {
    uint64_t collisions = 0;
    /* `indices_[ii] != h` compares the int32 entry with the unsigned hash (usual arithmetic conversions),
     * `indices_[ii] != -1` is an int comparison */
    while ((uint32_t)a->claims[ii] != h && a->claims[ii] != -1)
    {
      collisions++;
      if (++ii >= memory) ii = 0;
    }
    if (collisions > a->pressure) a->pressure = collisions;
    if (claim) a->claims[ii] = (int32_t)h;
  }
  return ii;
}

/* TileCodingProjector::_project(in, claim) (tile_coding.cpp:103-149) for this agent's projector */
static void agent_project(orc_agent *a, const double *in, int dims, int claim, uint32_t *hash, uint32_t *idx)
{
  uint32_t h[ORC_MAX_TILINGS];
  orc_project_hashes(a->c.resolution, a->c.wrapping, dims, a->c.tilings, in, h);
  for (int j = 0; j < a->c.tilings; ++j)
  {
    if (hash) hash[j] = h[j];
    idx[j] = feature_location(a, h[j], claim);
  }
}

uint64_t orc_pressure(const orc_agent *a) { return a->pressure; }

/* ---- QPolicy (q.cpp:94-107,129-155) ------------------------------------- */

static void q_values(orc_agent *a, const double *obs, uint32_t *hash, uint32_t *idx, double *q)
{
  const orc_config *c = &a->c;
  double in[ORC_MAX_DIMS];
  for (int i = 0; i < c->obs_dims; ++i) in[i] = obs[i];
  for (int aa = 0; aa < c->n_actions; ++aa)
  {
    in[c->obs_dims] = c->actions[aa];       /* extend(obs, variant) */
    /* Projector::project(base, variants, out): claims only with safe > 1 (tile_coding.h:66-72) */
    agent_project(a, in, c->obs_dims + 1, c->safe > 1, hash ? hash + aa * c->tilings : NULL, idx + aa * c->tilings);
    q[aa] = table_read(a->table, idx + aa * c->tilings, c->tilings, -DBL_MAX, DBL_MAX);
  }
}

/* ---- task start (pendulum.cpp:97-103, cart_pole.cpp:152-160, cart_double_pole.cpp:250-262) */
static double task_start(orc_agent *a, int test, double *state)
{
  const orc_config *c = &a->c;
  double r;
  if (c->rng == ORC_RNG_GLIBC) drand48_r(&a->randgen, &r);
  else
  {
    uint32_t w[4];
    philox_for(a, STREAM_START, a->episode_counter, w);
    r = u53(w[0], w[1]);
  }
  a->episode_counter++;
  orc_task_start(c, test, r, state);
  return r;
}

/* QPolicy::value (q.cpp:56-68) with GreedySampler::distribution (greedy.cpp:88-98) or
 * EpsilonGreedySampler::distribution (greedy.cpp:220-238, one epsilon, decay_ = 1): sum of q * dist in order. */
static double policy_value(const double *q, int A, double target_epsilon)
{
  int mai, man;
  double v = 0;
  findmax(q, A, &mai, &man);
  for (int ii = 0; ii < A; ++ii)
  {
    double d = (q[ii] == q[mai]) ? 1. / man : 0.;
    if (target_epsilon >= 0)
    {
      if (d == 1) d = 1 - 1. * target_epsilon;
      d += 1. * target_epsilon / A;
    }
    v += q[ii] * d;
  }
  return v;
}

/* ---- SARSAPredictor::criticize (sarsa.cpp:98-124), tau = 1; with algo = Q: QPredictor::criticize
 * (advantage.cpp:71-110, target = r + gamma max_a Q(s',a) by an fmax chain from -inf); with algo =
 * EXPECTED_SARSA: ExpectedSARSAPredictor::criticize (sarsa.cpp:167-194, target = r + gamma policy->value(s')).
 * q_next: Q(s', .) for every action (the predictor re-reads what the policy just read: no write in between). */
static void sarsa_update(orc_agent *a, double reward, const uint32_t *next_idx /* NULL if absorbing */,
                         const double *q_next, orc_record *rec)
{
  const orc_config *c = &a->c;
  const int T = c->tilings;
  double target = reward;
  if (next_idx)
  {
    double v;
    if (c->algo == ORC_ALGO_Q)
    {
      v = -INFINITY;
      for (int kk = 0; kk < c->n_actions; ++kk) v = fmax(v, q_next[kk]);
    }
    else if (c->algo == ORC_ALGO_EXPECTED_SARSA)
      v = policy_value(q_next, c->n_actions, c->target_epsilon);
    else
      v = table_read(a->table, next_idx, T, -DBL_MAX, DBL_MAX);
    target += c->gamma * v;                                                     /* pow(gamma,1) */
  }
  double qp = table_read(a->table, a->prev_idx, T, -DBL_MAX, DBL_MAX);
  double delta = target - qp;
  int swept = 0, entries = 0;

  table_write(a->table, a->prev_idx, T, target, a->alpha, -DBL_MAX, DBL_MAX);
  if (c->trace != ORC_TRACE_NONE)
  {
    double e = c->gamma * a->lambda;                                        /* pow(gamma*lambda,1) */
    trace_sweep(&a->trace, a->table, T, a->alpha * delta, e, -DBL_MAX, DBL_MAX, &swept, &entries);
    trace_add(&a->trace, c->trace, a->prev_idx, T, e);
  }
  a->swept_total += (uint64_t)swept;
  if (rec)
  {
    rec->q_prev = qp; rec->target = target; rec->delta = delta;
    rec->swept = swept; rec->swept_entries = entries;
  }
}

static void dump_trace(const orc_agent *a, orc_record *rec)
{
  const int T = a->c.tilings;
  int n = a->trace.len < ORC_MAX_TRACE ? a->trace.len : ORC_MAX_TRACE;
  rec->trace_len = a->trace.len;
  for (int k = 0; k < n; ++k)
  {
    int pos = a->trace.len - k - 1;
    rec->trace_decay[k] = a->trace.decay[pos];
    memcpy(&rec->trace_index[k * ORC_MAX_TILINGS], a->trace.idx[pos], (size_t)T * sizeof(uint32_t));
  }
}

/* ---- construction -------------------------------------------------------- */

orc_agent *orc_create(const orc_config *c, uint64_t agent_id)
{
  orc_agent *a = (orc_agent *)calloc(1, sizeof(orc_agent));
  if (!a) return NULL;
  a->c = *c;
  a->id = agent_id;
  a->S = orc_state_dims(c->dynamics);
  a->alpha = c->alpha; a->lambda = c->lambda; a->epsilon = c->epsilon;
  a->table = (double *)malloc((size_t)c->memory * sizeof(double));
  if (c->algo == ORC_ALGO_AC)
    a->actor = (double *)malloc((size_t)c->memory * sizeof(double));
  trace_clear(&a->trace);
  if (c->safe > 0)
  {
    a->claims = (int32_t *)malloc((size_t)c->memory * sizeof(int32_t));
    for (int32_t ii = 0; ii < c->memory; ++ii) a->claims[ii] = -1;
  }

  if (c->rng == ORC_RNG_GLIBC)
  {
    /* grld construction order for the shipped YAML (configurable.cpp:627-654 is
     * depth-first in document order): representation::configure -> reset ->
     * RandGen::instance() -> new Rand() -> srand48_r(lrand48()) (utils.h:90-93),
     * then sampler/epsilon_greedy configure -> new Rand(), then the test agent's
     * sampler/greedy -> new Rand(). */
    srand48_r(lrand48(), &a->randgen);
    if (a->actor)
    {
      /* AC YAML: the policy's (actor) representation is configured before the critic's; the action
       * policies draw from RandGen itself, no sampler objects exist */
      table_init(a, a->actor, STREAM_ACTOR_TABLE);
      table_init(a, a->table, STREAM_TABLE);
    }
    else
    {
      table_init(a, a->table, STREAM_TABLE);
      if (c->sampler != ORC_SAMPLER_SOFTMAX)            /* sampler/softmax owns no Rand (samplers/softmax.h) */
        srand48_r(lrand48(), &a->sampler);
      srand48_r(lrand48(), &a->test_sampler);
    }
  }
  else
  {
    table_init(a, a->table, STREAM_TABLE);
    if (a->actor) table_init(a, a->actor, STREAM_ACTOR_TABLE);
  }
  return a;
}

void orc_destroy(orc_agent *a)
{
  if (!a) return;
  free(a->table); free(a->actor); free(a->claims); free(a);
}

void orc_set_hyper(orc_agent *a, double alpha, double lambda, double epsilon)
{
  a->alpha = alpha; a->lambda = lambda; a->epsilon = epsilon;
}

double *orc_table(orc_agent *a) { return a->table; }
double *orc_actor_table(orc_agent *a) { return a->actor; }

void orc_stats(const orc_agent *a, uint64_t *steps, uint64_t *episodes, uint64_t *swept,
               double *last_return, double *last_test_return)
{
  if (steps) *steps = a->ss;
  if (episodes) *episodes = a->episodes;
  if (swept) *swept = a->swept_total;
  if (last_return) *last_return = a->last_return;
  if (last_test_return) *last_test_return = a->last_test_return;
}

void orc_get_state(const orc_agent *a, double *state, double *prev_obs, double *prev_action,
                   int32_t *in_episode)
{
  if (state) memcpy(state, a->state, sizeof(double) * ORC_MAX_STATE);
  if (prev_obs) memcpy(prev_obs, a->prev_obs, sizeof(double) * ORC_MAX_DIMS);
  if (prev_action) *prev_action = a->prev_action;
  if (in_episode) *in_episode = a->in_episode;
}

/* the action that the next environment step will apply */
double orc_pending_action(const orc_agent *a) { return a->action; }

static int ac_step(orc_agent *a, orc_record *rec);

/* safe > 0: the predictor's own projection of (prev_obs, prev_action), claiming (sarsa.cpp:102) */
static void reproject_prev(orc_agent *a)
{
  double in[ORC_MAX_DIMS];
  for (int i = 0; i < a->c.obs_dims; ++i) in[i] = a->prev_obs[i];
  in[a->c.obs_dims] = a->prev_action;
  agent_project(a, in, a->c.obs_dims + 1, 1, NULL, a->prev_idx);
}

/* ---- one agent-step (online_learning.cpp:154-222) ------------------------ */
int orc_step(orc_agent *a, orc_record *rec)
{
  const orc_config *c = &a->c;
  const int T = c->tilings, A = c->n_actions;
  uint32_t hash[ORC_MAX_ACTIONS * ORC_MAX_TILINGS], idx[ORC_MAX_ACTIONS * ORC_MAX_TILINGS];
  double q[ORC_MAX_ACTIONS], obs[ORC_MAX_DIMS], next[ORC_MAX_STATE], reward;
  int32_t terminal;
  act_draws d;

  if (c->algo == ORC_ALGO_AC) return ac_step(a, rec);
  if (rec) memset(rec, 0, sizeof(*rec));

  if (!a->in_episode)
  {
    /* online_learning.cpp:160: test episode every (test_interval+1)-th trial */
    a->test = (c->test_interval >= 0 &&
               a->tt % (uint64_t)(c->test_interval + 1) == (uint64_t)c->test_interval);
    double r = task_start(a, a->test, a->state);         /* modeled.cpp:132-158 */
    orc_observe(c, a->state, obs, &terminal);
    if (!a->test)
      trace_clear(&a->trace);                             /* TDAgent::start -> predictor finalize */
    q_values(a, obs, hash, idx, q);
    int ai = sample(a, q, a->test, &d);
    a->action = c->actions[ai];
    memcpy(a->prev_obs, obs, sizeof(obs));
    a->prev_action = a->action;
    memcpy(a->prev_idx, idx + ai * T, (size_t)T * sizeof(uint32_t));
    a->total_reward = 0;
    a->in_episode = 1;
    if (rec)
    {
      rec->started = 1; rec->start_draw = r;
      memcpy(rec->start_state, a->state, sizeof(double) * (size_t)a->S);
      rec->start_u = d.u; rec->start_r_explore = d.r_explore; rec->start_r_tie = d.r_tie;
      rec->start_action_index = ai;
      memcpy(rec->start_q, q, sizeof(double) * (size_t)A);
    }
  }

  orc_env_step(c, a->state, a->action, next, obs, &reward, &terminal);
  a->total_reward += reward;
  if (rec)
  {
    memcpy(rec->state, a->state, sizeof(double) * (size_t)a->S);
    rec->action = a->action;
    memcpy(rec->next, next, sizeof(double) * (size_t)a->S);
    memcpy(rec->obs, obs, sizeof(double) * (size_t)c->obs_dims);
    rec->reward = reward; rec->terminal = terminal; rec->test = a->test;
    rec->action_index = -1;
  }

  if (terminal == 2)
  {
    /* TDAgent::end (td.cpp:76-81): update without bootstrap; FixedAgent::end: nothing */
    if (!a->test)
    {
      if (c->safe > 0) reproject_prev(a);
      sarsa_update(a, reward, NULL, NULL, rec);
    }
  }
  else
  {
    /* TDAgent::step (td.cpp:63-74): act on the new observation, then update */
    q_values(a, obs, hash, idx, q);
    int ai = sample(a, q, a->test, &d);
    double action = c->actions[ai];
    if (rec)
    {
      rec->u = d.u; rec->r_explore = d.r_explore; rec->r_tie = d.r_tie; rec->explored = d.explored;
      for (int aa = 0; aa < A; ++aa)
      {
        memcpy(&rec->hash[aa * ORC_MAX_TILINGS], hash + aa * T, (size_t)T * sizeof(uint32_t));
        memcpy(&rec->index[aa * ORC_MAX_TILINGS], idx + aa * T, (size_t)T * sizeof(uint32_t));
        rec->q[aa] = q[aa];
      }
      rec->action_index = ai;
    }
    if (!a->test)
    {
      if (c->safe > 0)
      {
        /* SARSAPredictor::criticize (sarsa.cpp:102-108) projects (prev_obs, prev_action) and then (obs, action) itself,
         * through Projector::project(base, variant) -> project(in): BOTH claim their locations (tile_coding.h:62-65) */
        double in[ORC_MAX_DIMS];
        reproject_prev(a);
        for (int i = 0; i < c->obs_dims; ++i) in[i] = obs[i];
        in[c->obs_dims] = action;
        agent_project(a, in, c->obs_dims + 1, 1, NULL, idx + ai * T);
      }
      sarsa_update(a, reward, idx + ai * T, q, rec);
    }
    memcpy(a->prev_obs, obs, sizeof(obs));
    a->prev_action = action;
    memcpy(a->prev_idx, idx + ai * T, (size_t)T * sizeof(uint32_t));
    a->action = action;
  }
  if (rec) dump_trace(a, rec);

  memcpy(a->state, next, sizeof(double) * (size_t)a->S);
  int counted = !a->test;
  if (counted) a->ss++;
  if (terminal)
  {
    a->in_episode = 0;
    a->tt++;
    a->episodes++;
    if (a->test) a->last_test_return = a->total_reward;
    else a->last_return = a->total_reward;
    if (rec) { rec->episode_done = 1; rec->episode_return = a->total_reward; }
  }
  return counted;
}

uint64_t orc_run(orc_agent *a, uint64_t n_steps)
{
  uint64_t n = 0;
  while (n < n_steps)
    n += (uint64_t)orc_step(a, NULL);
  return n;
}

/* n agent-steps counting test-episode steps as well (what grlcu_run counts) */
void orc_run_total(orc_agent *a, uint64_t n_steps)
{
  for (uint64_t n = 0; n < n_steps; ++n) orc_step(a, NULL);
}

/* ---- actor-critic (cfg 4: cfg/cart_double_pole/ac.yaml) ----------------------------------
 * agent/td with policy/parameterized/action (action.cpp:99-158) and predictor/ac/action
 * (ac.cpp:72-110, update_method = proportional, no step limit) around predictor/critic/td
 * (predictors/td.cpp:68-91).  Actor and critic each own a projector and a table; the projectors
 * hash with the same function (tile_coding.h:78-151 has no per-object seed).
 * Record layout in AC mode: hash/index rows 0 = actor projection of obs', row 1 = critic projection
 * of obs'; q[0] = pi(s') (actor read, clamped), q[1] = V(s'), q[2] = u = pi(s) read inside update(),
 * q[3] = a' = clamp(pi(s') + n); noise / start_noise = the Gaussian term n. */

/* Rand::getNormal (utils.h:120-125): Box-Muller on two uniforms, U1 first */
static double get_normal(orc_agent *a, double mu, double sigma)
{
  double U1, U2;
  if (a->c.rng == ORC_RNG_GLIBC)
  {
    drand48_r(&a->randgen, &U1);
    drand48_r(&a->randgen, &U2);
  }
  else
  {
    uint32_t w[4];
    philox_for(a, STREAM_NOISE, a->act_counter++, w);
    U1 = u53(w[0], w[1]);
    U2 = u53(w[2], w[3]);
  }
  return sqrt(-2 * log(U1)) * cos(2 * M_PI * U2) * sigma + mu;
}

/* ActionPolicy::act(time, in, out) (action.cpp:129-158), one action dimension, theta = 1,
 * decay_ = 1 (decay_rate default 1).  test: the test agent's policy has sigma = 0. */
static double ac_act(orc_agent *a, const double *obs, int test, int time_is_zero,
                     uint32_t *hash, uint32_t *idx, double *pi, double *noise)
{
  const orc_config *c = &a->c;
  orc_project(c->actor_resolution, c->actor_wrapping, c->obs_dims, c->tilings, c->memory, obs, hash, idx);
  double out = table_read(a->actor, idx, c->tilings, c->action_min, c->action_max);   /* ActionPolicy::read */
  *pi = out;
  *noise = 0;
  if (time_is_zero) a->noise = 0.;                       /* n_ = ConstantVector(size, 0.) */
  if (!test && c->sigma)
  {
    a->noise = (1 - 1.) * a->noise + get_normal(a, 0., 1. * c->sigma);
    out += a->noise;
    *noise = a->noise;
  }
  return fmin(fmax(out, c->action_min), c->action_max);
}

/* ActionACPredictor::update (ac.cpp:72-110) around TDPredictor::criticize (td.cpp:68-91), tau = 1 */
static void ac_update(orc_agent *a, double reward, const uint32_t *next_cidx /* NULL if absorbing */, orc_record *rec)
{
  const orc_config *c = &a->c;
  const int T = c->tilings;
  /* ap = projector_->project(prev_obs); u = representation_->read(ap) */
  double u = table_read(a->actor, a->prev_aidx, T, c->action_min, c->action_max);
  /* critic */
  double target = reward;
  if (next_cidx)
    target += c->gamma * table_read(a->table, next_cidx, T, -DBL_MAX, DBL_MAX);
  double vp = table_read(a->table, a->prev_idx, T, -DBL_MAX, DBL_MAX);
  double delta = target - vp;
  int swept = 0, entries = 0;
  table_write(a->table, a->prev_idx, T, target, a->alpha, -DBL_MAX, DBL_MAX);
  if (c->trace != ORC_TRACE_NONE)
  {
    double e = c->gamma * a->lambda;
    trace_sweep(&a->trace, a->table, T, a->alpha * delta, e, -DBL_MAX, DBL_MAX, &swept, &entries);
    trace_add(&a->trace, c->trace, a->prev_idx, T, e);
  }
  /* actor: delta = critique * (prev_action - u); target_u = u + delta; write(ap, target_u, alpha) */
  double du = delta * (a->prev_action - u);
  double target_u = u + du;
  table_write(a->actor, a->prev_aidx, T, target_u, c->actor_alpha, c->action_min, c->action_max);
  a->swept_total += (uint64_t)swept;
  if (rec)
  {
    rec->q_prev = vp; rec->target = target; rec->delta = delta;
    rec->swept = swept; rec->swept_entries = entries; rec->q[2] = u;
  }
}

static int ac_step(orc_agent *a, orc_record *rec)
{
  const orc_config *c = &a->c;
  const int T = c->tilings;
  uint32_t ahash[ORC_MAX_TILINGS], aidx[ORC_MAX_TILINGS], chash[ORC_MAX_TILINGS], cidx[ORC_MAX_TILINGS];
  double obs[ORC_MAX_DIMS], next[ORC_MAX_STATE], reward, pi, noise;
  int32_t terminal;
  if (rec) memset(rec, 0, sizeof(*rec));

  if (!a->in_episode)
  {
    a->test = (c->test_interval >= 0 &&
               a->tt % (uint64_t)(c->test_interval + 1) == (uint64_t)c->test_interval);
    double r = task_start(a, a->test, a->state);
    orc_observe(c, a->state, obs, &terminal);
    /* TDAgent::start -> predictor_->finalize() is ActionACPredictor::finalize() here, which only calls the empty
     * Predictor::finalize() (ac.cpp:170-173, predictor.cpp:100-102): it does NOT reach the critic, so the critic's
     * eligibility trace is NOT cleared between episodes -- it keeps decaying across the episode boundary and the first
     * updates of an episode still move weights of the previous one.  (TDPredictor::finalize, td.cpp:93-99, runs only on
     * the experiment's `reset` at the start of a run.)  Found by compiling the reference's own ac.cpp / td.cpp against
     * this file (tests/test_oracle_vs_ref_agent.py); until then the oracle cleared the trace here. */
    a->action = ac_act(a, obs, a->test, 1, ahash, aidx, &pi, &noise);
    orc_project(c->resolution, c->wrapping, c->obs_dims, T, c->memory, obs, chash, cidx);
    memcpy(a->prev_obs, obs, sizeof(obs));
    a->prev_action = a->action;
    memcpy(a->prev_aidx, aidx, (size_t)T * sizeof(uint32_t));
    memcpy(a->prev_idx, cidx, (size_t)T * sizeof(uint32_t));
    a->total_reward = 0;
    a->in_episode = 1;
    if (rec)
    {
      rec->started = 1; rec->start_draw = r;
      memcpy(rec->start_state, a->state, sizeof(double) * (size_t)a->S);
      rec->start_noise = noise; rec->start_q[0] = pi; rec->start_q[3] = a->action;
    }
  }

  orc_env_step(c, a->state, a->action, next, obs, &reward, &terminal);
  a->total_reward += reward;
  if (rec)
  {
    memcpy(rec->state, a->state, sizeof(double) * (size_t)a->S);
    rec->action = a->action;
    memcpy(rec->next, next, sizeof(double) * (size_t)a->S);
    memcpy(rec->obs, obs, sizeof(double) * (size_t)c->obs_dims);
    rec->reward = reward; rec->terminal = terminal; rec->test = a->test;
    rec->action_index = -1;
  }

  if (terminal == 2)
  {
    if (!a->test) ac_update(a, reward, NULL, rec);
  }
  else
  {
    double action = ac_act(a, obs, a->test, 0, ahash, aidx, &pi, &noise);
    orc_project(c->resolution, c->wrapping, c->obs_dims, T, c->memory, obs, chash, cidx);
    if (rec)
    {
      memcpy(&rec->hash[0], ahash, (size_t)T * sizeof(uint32_t));
      memcpy(&rec->index[0], aidx, (size_t)T * sizeof(uint32_t));
      memcpy(&rec->hash[ORC_MAX_TILINGS], chash, (size_t)T * sizeof(uint32_t));
      memcpy(&rec->index[ORC_MAX_TILINGS], cidx, (size_t)T * sizeof(uint32_t));
      rec->q[0] = pi; rec->q[1] = table_read(a->table, cidx, T, -DBL_MAX, DBL_MAX);
      rec->q[3] = action; rec->noise = noise;
    }
    if (!a->test) ac_update(a, reward, cidx, rec);
    memcpy(a->prev_obs, obs, sizeof(obs));
    a->prev_action = action;
    memcpy(a->prev_aidx, aidx, (size_t)T * sizeof(uint32_t));
    memcpy(a->prev_idx, cidx, (size_t)T * sizeof(uint32_t));
    a->action = action;
  }
  if (rec) dump_trace(a, rec);

  memcpy(a->state, next, sizeof(double) * (size_t)a->S);
  int counted = !a->test;
  if (counted) a->ss++;
  if (terminal)
  {
    a->in_episode = 0;
    a->tt++;
    a->episodes++;
    if (a->test) a->last_test_return = a->total_reward;
    else a->last_return = a->total_reward;
    if (rec) { rec->episode_done = 1; rec->episode_return = a->total_reward; }
  }
  return counted;
}
