/* grlcu.h -- C ABI of the B200-native batched GRL learning loop (libgrlcu.so).
 *
 * This is the drop-in boundary for ONE hot path of wcaarls/grl: the on-line learning
 * loop (RK4 environment step -> tile-coding hash -> linear Q read -> epsilon-greedy ->
 * SARSA(lambda) update with enumerated trace), batched over many independent agents.
 * GRL itself has no C FFI: the path sits behind its C++ Configurable plugin API
 * (base/include/grl/configurable.h:704-826).  The callers of this ABI are the thin C++
 * Configurable shims in grl_b200/host/ (environment/model/cuda_batched,
 * projector/tile_coding/cuda, representation/linear/cuda, predictor/sarsa/cuda, and for the
 * sibling algorithms predictor/q/cuda, predictor/expected_sarsa/cuda, predictor/ac/action/cuda,
 * predictor/critic/td/cuda) and the Python ctypes harness used by tests/
 * and bench.py.
 *
 * Conventions: plain pointers and sizes, caller-owned HOST buffers unless a name says
 * "_dev", int return codes (0 = ok, non-zero = error, text in grlcu_last_error()),
 * opaque context handle, one context per host thread / GPU.  There is no CPU fallback:
 * every entry point fails with GRLCU_ERR_CUDA when no sm_100 device is usable.
 *
 * Each entry point cites the reference interface it replaces (paths relative to
 * /root/reference/base).
 */
#ifndef GRLCU_H_
#define GRLCU_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Array capacities of grlcu_config.  What the kernels are INSTANTIATED for is narrower and checked by grlcu_create
 * (GRLCU_ERR_UNSUPPORTED otherwise): tilings = 16; one action dimension with n_actions = 3 for every kernel (every
 * tile-coded configuration the reference ships for these three tasks: cfg/pendulum, cfg/cart_pole, cfg/cart_double_pole),
 * n_actions = 2, 4 or 5 on the uncached kernels (discretizer/uniform with other `steps`); obs_dims = 2 / 4 / 6 as the
 * dynamics dictate. */
#define GRLCU_MAX_DIMS 8
#define GRLCU_MAX_ACTIONS 8
#define GRLCU_MAX_STATE 7

enum { GRLCU_DYN_PENDULUM = 0, GRLCU_DYN_CART_POLE = 1, GRLCU_DYN_CART_DOUBLE_POLE = 2 };
enum { GRLCU_TRACE_NONE = 0, GRLCU_TRACE_REPLACING = 1, GRLCU_TRACE_ACCUMULATING = 2 };
/* predictor: predictor/critic/sarsa (sarsa.cpp:98-124) | predictor/ac/action + predictor/critic/td (ac.cpp, td.cpp) |
 * predictor/critic/q (advantage.cpp:71-110) | predictor/critic/expected_sarsa (sarsa.cpp:167-194) */
enum { GRLCU_ALGO_SARSA = 0, GRLCU_ALGO_AC = 1, GRLCU_ALGO_Q = 2, GRLCU_ALGO_EXPECTED_SARSA = 3 };
enum { GRLCU_SAMPLER_EPSILON_GREEDY = 0, GRLCU_SAMPLER_SOFTMAX = 1 };
enum { GRLCU_OK = 0, GRLCU_ERR_ARG = 1, GRLCU_ERR_CUDA = 2, GRLCU_ERR_UNSUPPORTED = 3,
       GRLCU_ERR_NOMEM = 4 };
enum { GRLCU_RESET_TABLES = 1, GRLCU_RESET_AGENTS = 2, GRLCU_RESET_STATS = 4 };

/* Parameters of the four GRL objects on the path, named after their YAML keys. */
typedef struct grlcu_config
{
  int32_t dynamics;           /* dynamics/pendulum | cart_pole | cart_double_pole */
  int32_t integration_steps;  /* model/dynamical integration_steps (modeled.cpp:234-240) */
  double  control_step;       /* model/dynamical control_step */
  double  timeout;            /* task timeout */
  double  randomization;      /* task randomization */
  int32_t tilings;            /* projector/tile_coding tilings (tile_coding.cpp:35-43); instantiated: 16 */
  int32_t memory;             /* projector/tile_coding memory = table entries PER AGENT */
  int32_t obs_dims;           /* task observation_dims */
  int32_t n_actions;          /* discretizer/uniform: number of discrete actions; instantiated: 2..5 (0 for GRLCU_ALGO_AC) */
  double  resolution[GRLCU_MAX_DIMS];   /* projector resolution (obs dims then action) */
  double  wrapping[GRLCU_MAX_DIMS];     /* projector wrapping */
  double  actions[GRLCU_MAX_ACTIONS];   /* discretizer values (uniform.cpp:60-94) */
  double  alpha, gamma, lambda, epsilon; /* predictor/critic/sarsa, sampler/epsilon_greedy */
  int32_t trace;              /* trace/enumerated/replacing | accumulating | none */
  int32_t test_interval;      /* experiment/online_learning test_interval (-1 = none) */
  double  init_min, init_max; /* representation/parameterized/linear init_min/init_max */
  uint64_t seed;              /* Philox key; stream = global agent id */
  int32_t rng;                /* ignored (oracle field) */
  int32_t algo;               /* GRLCU_ALGO_* */
  double  sigma;              /* policy/parameterized/action sigma (AC; action.cpp:40) */
  double  actor_alpha;        /* predictor/ac/action alpha (AC; ac.cpp:40) */
  double  action_min, action_max;       /* task action_min/max = actor output_min/max (AC) */
  double  actor_resolution[GRLCU_MAX_DIMS];  /* AC: the actor's projector; must equal resolution/wrapping, */
  double  actor_wrapping[GRLCU_MAX_DIMS];    /* which describe the critic's (obs dims only, no action)     */
  double  target_epsilon;     /* GRLCU_ALGO_EXPECTED_SARSA: the target policy's sampler (predictor `policy`,
                                 sarsa.cpp:145): < 0 sampler/greedy, else sampler/epsilon_greedy's epsilon */
  int32_t end_stop_penalty;   /* task/cart_pole/swingup end_stop_penalty (cart_pole.cpp:177,196; the reference's default
                                 is 1, cfg/cart_pole/ac_tc.yaml uses 0); ignored by the other tasks */
  int32_t sampler;            /* the learning policy's sampler: GRLCU_SAMPLER_EPSILON_GREEDY (greedy.cpp:144-218, uses
                                 `epsilon`) | GRLCU_SAMPLER_SOFTMAX (softmax.cpp:48-101, uses `softmax_tau`); test
                                 episodes always use sampler/greedy */
  double  softmax_tau;        /* sampler/softmax tau: temperature of the Boltzmann distribution (softmax.cpp:34) */
  int32_t safe;               /* projector/tile_coding safe (tile_coding.cpp:35-43, tile_coding.h:116-148): 0 = plain hashing;
                                 1 / 2 = collision-claim table (linear probing over int32 indices_[memory] per agent; claim on
                                 the predictor's projections / on every projection).  SARSA only; runs the uncached kernels */
} grlcu_config;

typedef struct grlcu_ctx grlcu_ctx;

/* Replay-mode inputs for one agent-step of every agent (host, SoA: [field][agent]).
 * Replaces nothing in the reference: it is the north-star's correctness harness
 * ("feeds identical states, actions and random draws to both implementations"). */
typedef struct grlcu_replay_in
{
  const int32_t  *started;       /* [n] 1: perform an episode start first */
  const double   *start_state;   /* [S][n] state after Task::start */
  const double   *start_u;       /* [n] draws of the start act() */
  const uint32_t *start_r_explore, *start_r_tie;
  const double   *state;         /* [S][n] state before the step (re-fed every step) */
  const double   *action;        /* [n] action to apply */
  const int32_t  *test;          /* [n] test episode flag */
  const double   *obs;           /* [obs_dims][n] observation fed to the agent half */
  const double   *reward;        /* [n] reward fed to the agent half */
  const int32_t  *terminal;      /* [n] terminal fed to the agent half */
  const double   *u;             /* [n] draws of this step's act() */
  const uint32_t *r_explore, *r_tie;
  const int32_t  *force_action;  /* [n] action index to commit (or NULL: use own choice) */
} grlcu_replay_in;

/* Replay-mode outputs (host, SoA).  Any pointer may be NULL. */
typedef struct grlcu_replay_out
{
  double   *next;          /* [S][n] RK4 result from in.state, in.action */
  double   *obs;           /* [obs_dims][n] observe(next) */
  double   *reward;        /* [n] evaluate(state, action, next) */
  int32_t  *terminal;      /* [n] */
  uint32_t *hash;          /* [A][T][n] tile hashes of (in.obs, a) */
  uint32_t *index;         /* [A][T][n] hash % memory */
  double   *q;             /* [A][n] */
  int32_t  *action_index;  /* [n] own choice (-1 when terminal == 2) */
  double   *start_q;       /* [A][n] (only meaningful where started) */
  int32_t  *start_action_index; /* [n] */
  double   *q_prev, *target, *delta;   /* [n] */
  int32_t  *swept;         /* [n] valid trace indices swept (V) */
  int32_t  *trace_len;     /* [n] after add() */
  uint32_t *trace_index;   /* [trace_cap][T][n] newest first, 0xFFFFFFFF = invalidated */
} grlcu_replay_out;

/* Per-agent results (host arrays of n_agents). Any pointer may be NULL. */
typedef struct grlcu_results
{
  double   *last_return;       /* return of the last finished learning episode */
  double   *last_test_return;  /* return of the last finished test episode */
  double   *return_sum;        /* sum of learning-episode returns */
  uint64_t *steps;             /* learning steps (ss, online_learning.cpp:219) */
  uint32_t *episodes;          /* finished episodes (learning + test) */
  uint64_t *swept;             /* sum of V over learning steps */
  uint64_t *table_writes;      /* random 8-byte table stores issued (roofline accounting) */
} grlcu_results;

const char *grlcu_last_error(void);
int grlcu_version(void);

/* Measurement aid (replaces nothing in the reference): the rates, in accesses per second, at which THIS device
 * serves independent random 8-byte reads (ld.global.L2::64B, what the table gathers use), writes and
 * read-modify-writes over a scratch buffer of `footprint_bytes` (a power of two >= 64 MiB, well beyond L2).
 * Every table access of the path is such an access, so these are the denominators of bench.py's
 * `roofline.random_access`.  Any output pointer may be NULL.  Takes ~50 ms. */
int grlcu_probe_random_access(int32_t device, uint64_t footprint_bytes, double *reads_per_s, double *writes_per_s,
                              double *rmws_per_s);

/* Creates the device-resident state of n_agents agents with global ids
 * [first_agent_id, first_agent_id + n_agents) on CUDA device `device`, launching on
 * `stream` (a cudaStream_t, NULL = default stream).  Tables are initialised like
 * LinearRepresentation::reset (linear.cpp:104-125) from Philox(seed, agent id).
 * Replaces: object construction + configure() of ModeledEnvironment/DynamicalModel
 * (modeled.cpp:59-121,234-252), TileCodingProjector (tile_coding.cpp:45-80),
 * LinearRepresentation (linear.cpp:66-102), SARSAPredictor (sarsa.cpp:48-59). */
int grlcu_create(const grlcu_config *cfg, int32_t n_agents, int64_t first_agent_id,
                 int32_t device, void *stream, grlcu_ctx **out);
int grlcu_destroy(grlcu_ctx *ctx);

/* reconfigure(action=reset) of the objects (linear.cpp:104-125, sarsa.cpp:61-67,
 * modeled.cpp:117-121).  what = OR of GRLCU_RESET_*. */
int grlcu_reset(grlcu_ctx *ctx, int32_t what);

/* Per-agent alpha / lambda / epsilon (cfg 5 hyper-parameter sweep).  NULL keeps the
 * config value.  Replaces per-instance YAML parameters of MultiExperiment / grlo. */
int grlcu_set_hyper(grlcu_ctx *ctx, const double *alpha, const double *lambda, const double *epsilon);

/* K2: TileCodingProjector::_project (tile_coding.cpp:103-149, tile_coding.h:78-151).
 * in: [n][dims] row-major, dims = obs_dims + 1.  hash/index: [n][tilings]. */
int grlcu_project(grlcu_ctx *ctx, int64_t n, const double *in, uint32_t *hash, uint32_t *index);

/* grlcu_project without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and results as grlcu_project (bit-identical), but the device staging of the inputs, hashes and indices
 * lives in one file-scope `static __device__` array of the library (768 KiB per GPU that loads it, held until exit)
 * instead of per-call cudaMalloc / cudaFree; the call performs no cudaMalloc, cudaFree or other allocating routine.
 * Limit: n <= GRLCU_NOALLOC_MAX_INPUTS, larger batches are rejected with GRLCU_ERR_ARG (split them, or use
 * grlcu_project).  Every call shares the array, which keeps the previous call's contents (nothing in it needs to start
 * at zero: every word that is read back is written first); calls that overlap on one GPU race on it, so the caller
 * serialises them. */
#define GRLCU_NOALLOC_MAX_INPUTS 4096
int grlcu_project_noalloc(grlcu_ctx *ctx, int64_t n, const double *in, uint32_t *hash, uint32_t *index);

/* K1: ModeledEnvironment::step body (modeled.cpp:160-213) = Task::actuate +
 * DynamicalModel::step (modeled.cpp:254-276) + Task::observe + Task::evaluate.
 * state/next: [n][S] row-major; action/reward: [n]; obs: [n][obs_dims]. */
int grlcu_model_step(grlcu_ctx *ctx, int64_t n, const double *state, const double *action,
                     double *next, double *obs, double *reward, int32_t *terminal);

/* Task::start + Task::observe for n independent episodes (pendulum.cpp:97-121, cart_pole.cpp:155-190):
 * u[n] are the uniform draws the reference takes from RandGen::get() (consumed even when the
 * randomization is 0).  state: [n][S] row-major; obs: [n][obs_dims]; terminal may be NULL. */
int grlcu_task_start(grlcu_ctx *ctx, int64_t n, int32_t test, const double *u, double *state, double *obs,
                     int32_t *terminal);

/* One-agent views of the representation, for the L2-signature shims:
 * LinearRepresentation::read (linear.cpp:136-184): value = (sum of table[idx[i]], in order) / n_idx;
 * LinearRepresentation::update(projection, delta) (linear.cpp:198-216): table[idx[i]] += delta for
 * every idx[i] != (uint64_t)-1, sequentially (duplicates accumulate). */
int grlcu_read(grlcu_ctx *ctx, int32_t agent, int32_t n_idx, const uint64_t *idx, double *value);
int grlcu_update(grlcu_ctx *ctx, int32_t agent, int32_t n_idx, const uint64_t *idx, double delta);
/* The same views of the ACTOR table of an actor-critic context (ActionPolicy's representation,
 * policies/action.cpp:99-113): the read is clamped to [action_min, action_max] (linear.cpp:179-180) and so is
 * every updated weight (limit = 1, linear.cpp:207-211). */
int grlcu_read_actor(grlcu_ctx *ctx, int32_t agent, int32_t n_idx, const uint64_t *idx, double *value);
int grlcu_update_actor(grlcu_ctx *ctx, int32_t agent, int32_t n_idx, const uint64_t *idx, double delta);

/* Fused K1-K4: n_steps iterations of online_learning.cpp:186-221 for every agent
 * (episode start/end handled on device).  Asynchronous on the context's stream.
 * cohort > 0 runs the agents in groups of `cohort` (each group does all n_steps before
 * the next starts) so a group's tables stay L2-resident; 0 = all agents at once. */
int grlcu_run(grlcu_ctx *ctx, int64_t n_steps, int32_t cohort);
int grlcu_sync(grlcu_ctx *ctx);

/* Plugin-facing end-to-end call with HOST buffers: uploads a batch of start states
 * ([n_agents][S] row-major; the next episode an agent BEGINS starts from its row, in
 * place of Task::start -- episodes in progress continue), runs n_steps agent-steps per
 * agent, downloads the results, and synchronises.  start_states may be NULL (Philox starts). */
int grlcu_run_host(grlcu_ctx *ctx, const double *start_states, int64_t n_steps, int32_t cohort,
                   grlcu_results *results);

/* One replayed agent-step for all agents (see grlcu_replay_in/out).
 * Actor-critic contexts (GRLCU_ALGO_AC) use the same structs: in.start_u / in.u carry the Gaussian
 * exploration term n of ActionPolicy::act (action.cpp:146-149) instead of a uniform, the integer draws and
 * force_action are ignored; out.hash / out.index hold ONE row of T (projection of in.obs; actor and
 * critic projectors share it), out.q four rows [pi(s'), V(s'), u = pi(s) read by the update, a'] and
 * out.start_q four rows [pi(s0), -, -, a0]; out.action_index / start_action_index are not written. */
int grlcu_replay_step(grlcu_ctx *ctx, const grlcu_replay_in *in, grlcu_replay_out *out);

int grlcu_get_results(grlcu_ctx *ctx, grlcu_results *results);

/* ParameterizedRepresentation::params()/setParams (representation.h:275-281) and the
 * raw-double checkpoint format of representation.h:201-263: `memory` doubles. */
int grlcu_get_table(grlcu_ctx *ctx, int32_t agent, double *out);
int grlcu_set_table(grlcu_ctx *ctx, int32_t agent, const double *in);

/* Actor table of an actor-critic context (GRLCU_ALGO_AC; policy/parameterized/action's
 * representation, policies/action.cpp:99-113): `memory` doubles, same checkpoint format. */
int grlcu_get_actor_table(grlcu_ctx *ctx, int32_t agent, double *out);
int grlcu_set_actor_table(grlcu_ctx *ctx, int32_t agent, const double *in);

/* Environment state ([S]) and pending action of one agent. */
int grlcu_get_state(grlcu_ctx *ctx, int32_t agent, double *state, double *action, int32_t *in_episode);

/* ---- several GPUs of one box ---------------------------------------------------------------------------------
 * One job = one context per GPU, each owning a contiguous range of global agent ids (the agents never exchange
 * anything while they learn).  A group drives the contexts together from one host thread (launches are
 * asynchronous, so the GPUs run concurrently) and performs the job's ONLY collective: an all-gather of the
 * per-agent results over NCCL (NVLink/NVSwitch), after which every GPU -- and the host, from the first context's
 * device -- holds the results of all agents in global id order.  libnccl.so.2 is loaded on first use; a group of
 * one context never touches NCCL.  All contexts must hold the same number of agents.
 * Replaces: MultiExperiment's role as "many independent runs at once" (experiments/multi.cpp:44-75: N instances on
 * N threads of one process) for the batched path. */
typedef struct grlcu_group grlcu_group;
int grlcu_device_count(void);                     /* usable CUDA devices (0 when there is none) */
int grlcu_group_create(grlcu_ctx *const *ctxs, int32_t n_ctx, grlcu_group **out);
int grlcu_group_destroy(grlcu_group *group);       /* does not destroy the contexts */
/* grlcu_run on every context (cohort 0). */
int grlcu_group_run(grlcu_group *group, int64_t n_steps);
/* Per-agent results of ALL agents of the group (host arrays of n_ctx * n_agents, context order = global id order).
 * collective_ms (may be NULL): device time of the all-gather alone, CUDA events on the first context's stream. */
int grlcu_group_results(grlcu_group *group, grlcu_results *results, double *collective_ms);

/* How the context stores its tables (measurement aid, replaces nothing in the reference): tile_major = 1 (slots from tile
 * coordinates) or 2 (visit-ordered from a pilot run) when the
 * tile-major physical layout is in use (the action words of one tile share a 32-byte sector and the
 * kernel computes the slot from the tile coordinates; logical indices, grlcu_get_table / grlcu_set_table and checkpoints
 * are unaffected), table_stride = doubles per agent, inbox_words = tile words of the task's observation box,
 * shared_words = those of them whose index collides with an earlier tile's (they live in that tile's sector).
 * Any pointer may be NULL. */
int grlcu_layout_info(grlcu_ctx *ctx, int32_t *tile_major, uint32_t *table_stride, uint32_t *inbox_words, uint32_t *shared_words);

/* projector/tile_coding safe > 0: per agent, the largest number of probe steps a lookup in its claim table has taken
 * (TileCodingProjector::pressure_, tile_coding.h:131-136: the reference warns above 7).  out: n_agents values. */
int grlcu_get_pressure(grlcu_ctx *ctx, uint32_t *out);

/* Trace capacity (entries) the context allocated per agent. */
int grlcu_trace_capacity(grlcu_ctx *ctx);
/* Counters for the roofline: launches of our kernels so far, and kernel name list. */
int64_t grlcu_launch_count(grlcu_ctx *ctx);
/* Device time (ms) spent in the fused run kernel since the last call (CUDA events on
 * the context's stream), and the number of run-kernel launches it covers. */
int grlcu_run_kernel_time(grlcu_ctx *ctx, double *ms, int64_t *launches);

#ifdef __cplusplus
}
#endif
#endif
