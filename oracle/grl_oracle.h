/* grl_oracle.h -- CPU restatement of GRL's inner on-line learning loop.
 *
 * TEST INFRASTRUCTURE ONLY.  This is the parity oracle for the B200 path:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product (grl_b200/, libgrlcu.so)
 * never links, imports or calls anything in oracle/.
 *
 * It restates, in plain C99 with no dependencies, the arithmetic of the
 * reference files below (paths relative to /root/reference/base), single
 * agent at a time, in the reference's own evaluation order.  Compile with
 * -O2 -ffp-contract=off and without -ffast-math / -march=native, which is
 * how the reference's CMakeLists.txt:19 builds (no FMA contraction).
 *
 *   src/environments/modeled.cpp:132-276      start/step, RK4
 *   src/environments/pendulum.cpp:40-145      dynamics + swing-up task
 *   src/environments/cart_pole.cpp:42-237     dynamics + swing-up task
 *   src/environments/cart_double_pole.cpp:46-127,250-310
 *   src/projectors/tile_coding.cpp:45-80,103-149, projectors/tile_coding.h:78-151
 *   include/grl/utils.h:70-137                safe_mod, Rand
 *   include/grl/projection.h:68-113           IndexProjection::ssub
 *   src/representations/linear.cpp:104-238    reset/read/write/update
 *   include/grl/representation.h:79-83        trace sweep
 *   include/grl/trace.h:135-263               enumerated traces
 *   src/samplers/greedy.cpp:47-86,144-218     greedy / epsilon-greedy
 *   src/discretizers/uniform.cpp:60-94,138-151
 *   src/policies/q.cpp:94-155, src/policies/action.cpp:99-158
 *   src/predictors/sarsa.cpp:98-132, td.cpp:68-99, ac.cpp:72-110
 *   src/agents/td.cpp:50-81, src/agents/fixed.cpp:48-62
 *   src/experiments/online_learning.cpp:154-222
 *
 * PARITY PIN STATUS (details: DESIGN.md section 3).  The reference needs Eigen 3, which is neither vendored nor installed in
 * the authoring container; everything else of it compiles there against a stand-in Eigen.  The oracle is pinned by
 *   (a) the reference's seeded learning curve tests/template/pendulum-sarsa-tc-0.txt, all 181 rows
 *       (tests/test_oracle_golden.py): RK4, pendulum task, hash, Q read, samplers, SARSA(lambda) write, replacing trace;
 *   (b) the reference's tests/template/cart_pole_balancing-pid-0.txt, all 10 rows (grl_oracle_pin.c,
 *       tests/test_oracle_vs_ref.py): cart-pole equations of motion + RK4 over 5 states;
 *   (c) the reference's OWN environment sources (pendulum.cpp, cart_pole.cpp, cart_double_pole.cpp) compiled from
 *       /root/reference against stand-in headers into oracle/_ref/libgrlref_env.so: eom / start / actuate / observe /
 *       evaluate bit-exact on 1e5 random inputs per function (tests/test_oracle_vs_ref.py);
 *   (d) an independent byte-wise MurmurHash2 for the tile hash, Random123's known answers for Philox;
 *   (e) the reference's OWN agent-side sources (tile_coding.cpp, linear.cpp, enumerated.cpp, uniform.cpp, greedy.cpp, q.cpp,
 *       action.cpp, predictor.cpp, sarsa.cpp, td.cpp, ac.cpp, agents/td.cpp, agents/fixed.cpp) compiled from /root/reference with
 *       the reference's own headers into oracle/_ref/libgrlref_agent.so (stand-ins: Eigen, one YAML type) and fed this oracle's
 *       observations: every action and the final weight tables bit-exact over thousands of steps for the SARSA agent
 *       (pendulum, cart-pole; replacing / accumulating / no trace; test episodes; hash collisions) and for the actor-critic
 *       agent on all three dynamics; likewise Q-learning, expected SARSA, sampler/softmax, the safe = 1|2 claim table, and
 *       DynamicalModel::step of the reference's own modeled.cpp for the 3-, 5- and 7-state systems against orc_rk4
 *       (tests/test_oracle_vs_ref_agent.py); and the closed loop of the reference's own tasks, model and agents (refag_run)
 *       against orc_step free-running from the same seed: states, counters and tables the same bits.  That run corrected
 *       ac_step: the reference does not clear the critic's trace between episodes.
 *   (f) grld itself built from the reference's translation units (deployer, YAML configurator, experiment loop, modeled
 *       environment, ...; vendored yaml-cpp; stand-in Eigen) into oracle/_ref/grld_ref: it writes the reference's golden curve
 *       of (a), and the rows it writes for 12 configurations equal this oracle's (tests/test_oracle_vs_grld_ref.py).
 * PARITY UNPINNED: the summation order inside Eigen's Matrix3d::inverse() and 3x3 products (cart-double-pole; the
 * stand-in shares the oracle's assumption).  Nothing else.
 */
#ifndef GRL_ORACLE_H_
#define GRL_ORACLE_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_DIMS 8      /* projector input dims (obs + action) */
#define ORC_MAX_ACTIONS 8
#define ORC_MAX_TILINGS 32
#define ORC_MAX_STATE 7
#define ORC_MAX_TRACE 64    /* trace entries dumped per record */
#define ORC_TRACE_STORE 160 /* trace entries an agent can hold */

enum { ORC_DYN_PENDULUM = 0, ORC_DYN_CART_POLE = 1, ORC_DYN_CART_DOUBLE_POLE = 2 };
enum { ORC_TRACE_NONE = 0, ORC_TRACE_REPLACING = 1, ORC_TRACE_ACCUMULATING = 2 };
enum { ORC_ALGO_SARSA = 0, ORC_ALGO_AC = 1, ORC_ALGO_Q = 2, ORC_ALGO_EXPECTED_SARSA = 3 };
enum { ORC_RNG_PHILOX = 0, ORC_RNG_GLIBC = 1 };
enum { ORC_SAMPLER_EPSILON_GREEDY = 0, ORC_SAMPLER_SOFTMAX = 1 };

/* Same field order as grlcu_config (include/grlcu.h) so one ctypes Structure
 * serves both; the two libraries share no code. */
typedef struct orc_config
{
  int32_t dynamics;           /* ORC_DYN_* */
  int32_t integration_steps;  /* model/dynamical: integration_steps (5) */
  double  control_step;       /* model/dynamical: control_step */
  double  timeout;            /* task: timeout */
  double  randomization;      /* task: randomization */
  int32_t tilings;            /* projector/tile_coding: tilings */
  int32_t memory;             /* projector/tile_coding: memory (entries per agent) */
  int32_t obs_dims;           /* 2 / 4 / 6 */
  int32_t n_actions;          /* discretizer/uniform: prod(steps) (SARSA) */
  double  resolution[ORC_MAX_DIMS];
  double  wrapping[ORC_MAX_DIMS];
  double  actions[ORC_MAX_ACTIONS];   /* discretizer values (1-D action) */
  double  alpha, gamma, lambda, epsilon;
  int32_t trace;              /* ORC_TRACE_* */
  int32_t test_interval;      /* experiment: test_interval (-1 = no test episodes) */
  double  init_min, init_max; /* representation: init_min/init_max */
  uint64_t seed;
  int32_t rng;                /* ORC_RNG_* (oracle only; ignored by grlcu) */
  int32_t algo;               /* ORC_ALGO_* */
  /* actor-critic (cfg 4) */
  double  sigma;              /* policy/action: sigma */
  double  actor_alpha;        /* predictor/ac/action: alpha */
  double  action_min, action_max;
  double  actor_resolution[ORC_MAX_DIMS];
  double  actor_wrapping[ORC_MAX_DIMS];
  /* predictor/critic/expected_sarsa: the target policy's sampler; < 0: sampler/greedy, else epsilon-greedy */
  double  target_epsilon;
  int32_t end_stop_penalty;   /* task/cart_pole/swingup: end_stop_penalty (cart_pole.h:75 default 1) */
  int32_t sampler;            /* ORC_SAMPLER_*: sampler/epsilon_greedy | sampler/softmax (learning policy) */
  double  softmax_tau;        /* sampler/softmax: tau */
  int32_t safe;               /* projector/tile_coding: safe (0 off, 1 claim on write, 2 claim always; tile_coding.h:116-148) */
} orc_config;

/* Everything observable about one agent-step (one iteration of
 * online_learning.cpp:186-221), for replay-mode comparison. */
typedef struct orc_record
{
  /* episode start performed at the top of this step (0/1) and what it drew */
  int32_t  started;
  double   start_draw;                /* task start uniform (always consumed) */
  double   start_state[ORC_MAX_STATE];
  double   start_u;                   /* sampler draws of the start act() */
  uint32_t start_r_explore, start_r_tie;
  int32_t  start_action_index;
  double   start_q[ORC_MAX_ACTIONS];
  double   noise, start_noise;        /* AC: Gaussian exploration term n */
  double   state[ORC_MAX_STATE];      /* state before the step */
  double   action;                    /* action applied */
  double   next[ORC_MAX_STATE];       /* state after RK4 */
  double   obs[ORC_MAX_DIMS];         /* observation of next */
  double   reward;
  int32_t  terminal;                  /* 0 / 1 / 2 */
  int32_t  test;                      /* test episode (no learning) */
  /* draws consumed by the sampler in this step's act() */
  double   u;                         /* uniform in [0,1) */
  uint32_t r_explore;                 /* raw integer used as r %% A when exploring */
  uint32_t r_tie;                     /* raw integer used as r %% man on ties */
  int32_t  explored;                  /* 1 if exploratory */
  /* act(): projections of (obs, a) for all actions */
  uint32_t hash[ORC_MAX_ACTIONS * ORC_MAX_TILINGS];
  uint32_t index[ORC_MAX_ACTIONS * ORC_MAX_TILINGS];
  double   q[ORC_MAX_ACTIONS];
  int32_t  action_index;              /* chosen a' (-1 if absorbing: no act) */
  /* update */
  double   q_prev;                    /* read(p) */
  double   target;
  double   delta;
  int32_t  swept;                     /* V: valid trace indices swept */
  int32_t  swept_entries;             /* trace entries visited by the sweep */
  /* trace after add(): newest first */
  int32_t  trace_len;
  double   trace_decay[ORC_MAX_TRACE];
  uint32_t trace_index[ORC_MAX_TRACE * ORC_MAX_TILINGS]; /* 0xFFFFFFFF = invalidated */
  /* episode bookkeeping after the step */
  int32_t  episode_done;
  double   episode_return;
} orc_record;

typedef struct orc_agent orc_agent;

/* ---- primitives (stateless) ---- */
uint32_t orc_hash(const int32_t *ints, uint32_t n);
/* independent formulation: byte-wise MurmurHash2 (Appleby) with adjusted seed */
uint32_t orc_murmur2_bytes(const void *key, int len, uint32_t seed);
void orc_scaling(const double *resolution, const double *wrapping, int dims, int tilings,
                 double *scaling, int32_t *iwrap);
void orc_project(const double *resolution, const double *wrapping, int dims, int tilings,
                 int32_t memory, const double *in, uint32_t *hash, uint32_t *index);
/* the tile hashes alone (createHashSum per tiling), before getFeatureLocation */
void orc_project_hashes(const double *resolution, const double *wrapping, int dims, int tilings,
                        const double *in, uint32_t *hash);
void orc_eom(int dynamics, const double *x, double u, double *xd);
void orc_rk4(int dynamics, double tau, int steps, const double *x, double u, double *next);
int  orc_state_dims(int dynamics);
void orc_observe(const orc_config *c, const double *x, double *obs, int32_t *terminal);
double orc_evaluate(const orc_config *c, const double *x, double action, const double *next);
void orc_task_start(const orc_config *c, int test, double r, double *state);
/* pin against tests/template/cart_pole_balancing-pid-0.txt (grl_oracle_pin.c) */
void orc_pin_cart_pole_pid(long seed, const double *p_gains, double timeout, int trials, double *rewards, int32_t *steps);
/* one environment step: actuate + rk4 + observe + evaluate (modeled.cpp:160-213) */
void orc_env_step(const orc_config *c, const double *x, double action,
                  double *next, double *obs, double *reward, int32_t *terminal);

/* Philox4x32-10 (Salmon et al. 2011), the framework's free-running RNG */
void orc_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                uint32_t k0, uint32_t k1, uint32_t out[4]);

/* ---- agents ---- */
orc_agent *orc_create(const orc_config *c, uint64_t agent_id);
void orc_destroy(orc_agent *a);
/* per-agent hyper-parameter overrides (cfg 5 sweep) */
void orc_set_hyper(orc_agent *a, double alpha, double lambda, double epsilon);
/* glibc mode only: process-global srand48(seed), then construction draws in
 * reference order (representation RandGen, epsilon sampler, greedy sampler). */
void orc_glibc_seed(long seed);
double *orc_table(orc_agent *a);            /* memory doubles (critic / Q) */
double *orc_actor_table(orc_agent *a);      /* memory doubles (AC only) */
/* advance one agent-step; rec may be NULL.  Handles episode start / end
 * internally exactly like online_learning.cpp:154-222.  Returns 1 if the step
 * counted as a learning step (ss++), 0 for a test-episode step. */
int orc_step(orc_agent *a, orc_record *rec);
/* run n learning steps without recording (CPU baseline timing) */
uint64_t orc_run(orc_agent *a, uint64_t n_steps);
/* projector/tile_coding safe > 0: the largest number of probe steps any lookup of this agent took (pressure_, tile_coding.h:131-136) */
uint64_t orc_pressure(const orc_agent *a);
/* run n agent-steps, test-episode steps included (the count grlcu_run takes) */
void orc_run_total(orc_agent *a, uint64_t n_steps);
/* statistics */
void orc_stats(const orc_agent *a, uint64_t *steps, uint64_t *episodes, uint64_t *swept,
               double *last_return, double *last_test_return);
double orc_pending_action(const orc_agent *a);
/* current state (for replay feeding) */
void orc_get_state(const orc_agent *a, double *state, double *prev_obs, double *prev_action,
                   int32_t *in_episode);

#ifdef __cplusplus
}
#endif
#endif
