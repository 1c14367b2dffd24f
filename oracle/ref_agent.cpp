/* oracle/ref_agent.cpp -- harness around the reference's OWN agent-side classes.  TEST INFRASTRUCTURE ONLY.
 *
 * oracle/Makefile (target `ref`) compiles, from where they lie under /root/reference/base/src,
 *   projectors/tile_coding.cpp  representations/linear.cpp  traces/enumerated.cpp  discretizers/uniform.cpp
 *   samplers/greedy.cpp  policies/q.cpp  policies/action.cpp  predictor.cpp  predictors/sarsa.cpp  predictors/td.cpp
 *   predictors/ac.cpp  agents/td.cpp  agents/fixed.cpp  environments/modeled.cpp  environments/{pendulum,cart_pole,cart_double_pole}.cpp
 * with the reference's own headers (vector.h, configurable.h, configuration.h, utils.h, trace.h, ...) and links them with
 * this file into oracle/_ref/libgrlref_agent.so.  The only stand-ins are Eigen (oracle/ref_shim_eigen/Eigen/Dense: the
 * library is absent here) and the YAML node type one declaration names; this file supplies the handful of globals that
 * live in the reference's configurable.cpp / grl.cpp (YAML-dependent, not compiled) and wires the objects up the way
 * grld does for cfg/pendulum/sarsa_tc.yaml (agent/td + policy/discrete/q + predictor/critic/sarsa, test agent
 * agent/fixed + sampler/greedy) and cfg/cart_double_pole/ac.yaml (agent/td + policy/parameterized/action +
 * predictor/ac/action around predictor/critic/td), in grld's construction order, because construction order decides
 * which drand48 stream each object gets (utils.h:84-137).
 *
 * tests/test_oracle_vs_ref_agent.py feeds the observations, rewards and terminal flags of an oracle run (glibc RNG mode,
 * same srand48 seed) to these objects and compares the actions step by step and the weight tables at the end, bit for bit.
 */
#include <stdint.h>
#include <string.h>
#include <stdlib.h>

#include <grl/configurable.h>
#include <grl/projectors/tile_coding.h>
#include <grl/representations/linear.h>
#include <grl/trace.h>
#include <grl/discretizers/uniform.h>
#include <grl/samplers/greedy.h>
#include <grl/samplers/softmax.h>
#include <grl/policies/q.h>
#include <grl/policies/action.h>
#include <grl/predictors/sarsa.h>
#include <grl/predictors/advantage.h>
#include <grl/predictors/td.h>
#include <grl/predictors/ac.h>
#include <grl/agents/td.h>
#include <grl/agents/fixed.h>
#include <grl/environments/pendulum.h>
#include <grl/environments/cart_pole.h>
#include <grl/environments/cart_double_pole.h>

extern "C" {
#include "grl_oracle.h"
}

using namespace grl;

/* ---- globals of the reference's configurable.cpp / grl.cpp (not compiled: they need yaml-cpp) ------------------------ */
namespace grl {
unsigned char grl_log_verbosity__ = 0;                       /* errors only */
const char *grl_log_levels__[] = { "(ERR)", "(WRN)", "(NTC)", "(INF)", "(DBG)", "(CRL)", "(TIM)" };
pthread_once_t RandGen::once_ = PTHREAD_ONCE_INIT;
pthread_mutex_t RandGen::mutex_;
pthread_key_t RandGen::key_;
DEFINE_FACTORY(Configurable)                                  /* the reference's own macro (factory.h:52-75), as configurable.cpp uses it */
DECLARE_TYPE_NAME(double)
DECLARE_TYPE_NAME(int)
DECLARE_TYPE_NAME(size_t)
DECLARE_TYPE_NAME(LargeVector)
}

namespace {

Vector vec(const double *d, int n) { Vector v(n); for (int i = 0; i < n; ++i) v[i] = d[i]; return v; }

struct RefAgent
{
  orc_config c;
  bool ac;
  /* SARSA */
  TileCodingProjector proj;
  LinearRepresentation repr;
  Trace *trace;
  UniformDiscretizer disc;
  EpsilonGreedySampler sampler;
  GreedySampler test_sampler;
  SoftmaxSampler softmax;
  GreedySampler target_greedy;                 /* expected SARSA: the target policy's sampler */
  EpsilonGreedySampler target_egreedy;
  QPolicy policy, test_policy, target_policy;
  SARSAPredictor sarsa;
  QPredictor qlearning;
  ExpectedSARSAPredictor expected_sarsa;
  /* actor-critic */
  TileCodingProjector actor_proj;
  LinearRepresentation actor_repr;
  ActionPolicy action_policy, test_action_policy;
  TDPredictor critic;
  ActionACPredictor acp;
  /* agents */
  TDAgent agent;
  FixedAgent test_agent;
  /* environment pieces for the closed loop (refag_run) */
  PendulumDynamics pd; CartPoleDynamics cpd; CartDoublePoleDynamics cdpd;
  PendulumSwingupTask pt; CartPoleSwingupTask cpt; CartDoublePoleBalancingTask cdpt;
  DynamicalModel model;
  Task *task;
  Vector state;
  bool in_episode; int test; uint64_t tt, ss; double total_reward;
  Action action;
  RefAgent() : trace(NULL), task(NULL), in_episode(false), test(0), tt(0), ss(0), total_reward(0) { }
  ~RefAgent() { delete trace; }
};

void configure_projector(TileCodingProjector &p, const orc_config &c, const double *res, const double *wrap, int dims)
{
  Configuration cfg;
  cfg.set("tilings", c.tilings);
  cfg.set("memory", c.memory);
  cfg.set("safe", c.safe);
  cfg.set("resolution", vec(res, dims));
  cfg.set("wrapping", vec(wrap, dims));
  p.configure(cfg);
}

void configure_representation(LinearRepresentation &r, const orc_config &c, bool clamp)
{
  Configuration cfg;
  cfg.set("interval", 0);
  cfg.set("tau", 1.0);
  cfg.set("file_suffix", "");
  cfg.set("memory", c.memory);
  cfg.set("outputs", 1);
  cfg.set("limit", 1);
  cfg.set("init_min", VectorConstructor(c.init_min));
  cfg.set("init_max", VectorConstructor(c.init_max));
  if (clamp)
  {
    cfg.set("output_min", VectorConstructor(c.action_min));
    cfg.set("output_max", VectorConstructor(c.action_max));
  }
  else
  {
    cfg.set("output_min", Vector());
    cfg.set("output_max", Vector());
  }
  r.configure(cfg);
  /* configure() ends in Configurable::reset() (linear.cpp:100-101), which walks the object's subtree through its
   * ObjectConfigurator (configurable.h:764-776).  There is no configurator tree here (it is the YAML side of the
   * reference), and the representation is a leaf: deliver the same message to the object directly -> memory draws from
   * the thread's RandGen stream (linear.cpp:108-125) */
  Configuration reset;
  reset.set("action", "reset");
  r.reconfigure(reset);
}

Trace *make_trace(const orc_config &c)
{
  Trace *t = NULL;
  if (c.trace == ORC_TRACE_REPLACING) t = new ReplacingEnumeratedTrace();
  else if (c.trace == ORC_TRACE_ACCUMULATING) t = new AccumulatingEnumeratedTrace();
  if (t) { Configuration cfg; t->configure(cfg); }
  return t;
}

template <class P>
void configure_critic(P &p, const orc_config &c, Projector *proj, Representation *repr, Trace *trace)
{
  Configuration cfg;
  cfg.set("importer", ""); cfg.set("exporter", "");
  cfg.set("projector", (void *)static_cast<Configurable *>(proj));
  cfg.set("representation", (void *)static_cast<Configurable *>(repr));
  if (trace) cfg.set("trace", (void *)static_cast<Configurable *>(trace)); else cfg.set("trace", "");
  cfg.set("alpha", c.alpha); cfg.set("gamma", c.gamma); cfg.set("lambda", c.lambda);
  p.configure(cfg);
}

void configure_action_policy(ActionPolicy &p, const orc_config &c, Projector *proj, Representation *repr, double sigma)
{
  Configuration cfg;
  cfg.set("projector", (void *)static_cast<Configurable *>(proj));
  cfg.set("representation", (void *)static_cast<Configurable *>(repr));
  cfg.set("renormalize", 0);
  cfg.set("sigma", VectorConstructor(sigma));
  cfg.set("theta", Vector());
  cfg.set("decay_rate", 1.0);
  cfg.set("decay_min", 0.0);
  cfg.set("output_min", VectorConstructor(c.action_min));
  cfg.set("output_max", VectorConstructor(c.action_max));
  p.configure(cfg);
}

void configure_q_policy(QPolicy &p, Discretizer *d, Projector *proj, Representation *repr, Sampler *s)
{
  Configuration cfg;
  cfg.set("discretizer", (void *)static_cast<Configurable *>(d));
  cfg.set("projector", (void *)static_cast<Configurable *>(proj));
  cfg.set("representation", (void *)static_cast<Configurable *>(repr));
  cfg.set("sampler", (void *)static_cast<Configurable *>(s));
  p.configure(cfg);
}

}  // namespace

extern "C" {

/* Creates the calling thread's RandGen instance (utils.h:160-175: `new Rand()` on first use, which consumes one
 * lrand48()).  Call once BEFORE seeding, so that every refag_create below starts from the same place. */
void refag_prepare(void) { (void)RandGen::instance(); }

/* The caller has done refag_prepare() once; srand48(seed) happens inside (see below).  In grld the thread's RandGen is
 * born inside the first representation's reset(): Rand() does srand48_r(lrand48()) (utils.h:90-93).  The instance of this
 * thread already exists, so it is re-seeded here the way a new one would be -- same call, same place in the lrand48()
 * sequence (nothing before the first representation draws).  Returns NULL and prints the reference's exception on failure. */
void *refag_create(const orc_config *c, long seed)
{
  RefAgent *a = new RefAgent();
  a->c = *c;
  a->ac = c->algo == ORC_ALGO_AC;
  try
  {
    /* expected SARSA's target policy owns a sampler whose Rand() consumes an lrand48() at construction but which never
     * draws (ValuePolicy::value uses its distribution(), q.cpp).  The oracle does not model that object's stream, so it
     * is built before the seed: the seeded sequence then starts where the oracle's does. */
    Sampler *target_sampler = NULL;
    if (c->algo == ORC_ALGO_EXPECTED_SARSA)
    {
      Configuration cfg;
      cfg.set("epsilon", VectorConstructor(c->target_epsilon < 0 ? 0. : c->target_epsilon));
      cfg.set("decay_rate", 1.0);
      cfg.set("decay_min", 0.0);
      if (c->target_epsilon < 0) { a->target_greedy.configure(cfg); target_sampler = &a->target_greedy; }
      else { a->target_egreedy.configure(cfg); target_sampler = &a->target_egreedy; }
    }
    srand48(seed);
    RandGen::instance()->init(lrand48());
    if (!a->ac)
    {
      const int dims = c->obs_dims + 1;
      configure_projector(a->proj, *c, c->resolution, c->wrapping, dims);
      configure_representation(a->repr, *c, false);          /* the table, from the RandGen stream seeded above (lrand48() #1) */
      a->trace = make_trace(*c);
      {
        Configuration cfg;
        cfg.set("min", VectorConstructor(c->actions[0]));
        cfg.set("max", VectorConstructor(c->actions[c->n_actions - 1]));
        cfg.set("steps", VectorConstructor((double)c->n_actions));
        a->disc.configure(cfg);
      }
      Sampler *learning_sampler;
      if (c->sampler == ORC_SAMPLER_SOFTMAX)
      {
        Configuration cfg;
        cfg.set("tau", c->softmax_tau);
        a->softmax.configure(cfg);                              /* owns no Rand: draws from the thread's RandGen */
        learning_sampler = &a->softmax;
      }
      else
      {
        Configuration cfg;
        cfg.set("epsilon", VectorConstructor(c->epsilon));
        cfg.set("decay_rate", 1.0);
        cfg.set("decay_min", 0.0);
        a->sampler.configure(cfg);                              /* new Rand(): lrand48() #2 */
        learning_sampler = &a->sampler;
      }
      configure_q_policy(a->policy, &a->disc, &a->proj, &a->repr, learning_sampler);
      Predictor *predictor;
      if (c->algo == ORC_ALGO_Q)
      {
        Configuration cfg;
        cfg.set("importer", ""); cfg.set("exporter", "");
        cfg.set("discretizer", (void *)static_cast<Configurable *>(&a->disc));
        cfg.set("projector", (void *)static_cast<Configurable *>(&a->proj));
        cfg.set("representation", (void *)static_cast<Configurable *>(&a->repr));
        if (a->trace) cfg.set("trace", (void *)static_cast<Configurable *>(a->trace)); else cfg.set("trace", "");
        cfg.set("alpha", c->alpha); cfg.set("gamma", c->gamma); cfg.set("lambda", c->lambda);
        a->qlearning.configure(cfg);
        predictor = &a->qlearning;
      }
      else if (c->algo == ORC_ALGO_EXPECTED_SARSA)
      {
        configure_q_policy(a->target_policy, &a->disc, &a->proj, &a->repr, target_sampler);
        Configuration cfg;
        cfg.set("importer", ""); cfg.set("exporter", "");
        cfg.set("projector", (void *)static_cast<Configurable *>(&a->proj));
        cfg.set("representation", (void *)static_cast<Configurable *>(&a->repr));
        cfg.set("policy", (void *)static_cast<Configurable *>(&a->target_policy));
        if (a->trace) cfg.set("trace", (void *)static_cast<Configurable *>(a->trace)); else cfg.set("trace", "");
        cfg.set("alpha", c->alpha); cfg.set("gamma", c->gamma); cfg.set("lambda", c->lambda);
        a->expected_sarsa.configure(cfg);
        predictor = &a->expected_sarsa;
      }
      else
      {
        configure_critic(a->sarsa, *c, &a->proj, &a->repr, a->trace);
        predictor = &a->sarsa;
      }
      {
        Configuration cfg;
        cfg.set("policy", (void *)static_cast<Configurable *>(&a->policy));
        cfg.set("predictor", (void *)static_cast<Configurable *>(predictor));
        a->agent.configure(cfg);
      }
      { Configuration cfg; a->test_sampler.configure(cfg); }  /* new Rand(): lrand48() #3 */
      configure_q_policy(a->test_policy, &a->disc, &a->proj, &a->repr, &a->test_sampler);
      { Configuration cfg; cfg.set("policy", (void *)static_cast<Configurable *>(&a->test_policy)); a->test_agent.configure(cfg); }
    }
    else
    {
      /* cfg/cart_double_pole/ac.yaml: the policy (actor projector + representation) is configured before the predictor's
       * critic; the action policies own no Rand, they draw from the thread's RandGen */
      configure_projector(a->actor_proj, *c, c->actor_resolution, c->actor_wrapping, c->obs_dims);
      configure_representation(a->actor_repr, *c, true);     /* the actor table, from the RandGen stream seeded above */
      configure_action_policy(a->action_policy, *c, &a->actor_proj, &a->actor_repr, c->sigma);
      configure_projector(a->proj, *c, c->resolution, c->wrapping, c->obs_dims);
      configure_representation(a->repr, *c, false);           /* the critic table, same stream */
      a->trace = make_trace(*c);
      configure_critic(a->critic, *c, &a->proj, &a->repr, a->trace);
      {
        Configuration cfg;
        cfg.set("importer", ""); cfg.set("exporter", "");
        cfg.set("projector", (void *)static_cast<Configurable *>(&a->actor_proj));
        cfg.set("representation", (void *)static_cast<Configurable *>(&a->actor_repr));
        cfg.set("critic", (void *)static_cast<Configurable *>(&a->critic));
        cfg.set("alpha", c->actor_alpha);
        cfg.set("update_method", "proportional");
        cfg.set("step_limit", Vector());
        a->acp.configure(cfg);
      }
      {
        Configuration cfg;
        cfg.set("policy", (void *)static_cast<Configurable *>(&a->action_policy));
        cfg.set("predictor", (void *)static_cast<Configurable *>(&a->acp));
        a->agent.configure(cfg);
      }
      configure_action_policy(a->test_action_policy, *c, &a->actor_proj, &a->actor_repr, 0.);
      { Configuration cfg; cfg.set("policy", (void *)static_cast<Configurable *>(&a->test_action_policy)); a->test_agent.configure(cfg); }
    }
  }
  catch (std::exception &e)
  {
    fprintf(stderr, "refag_create: %s\n", e.what());
    delete a;
    return NULL;
  }
  return a;
}

void refag_destroy(void *h) { delete (RefAgent *)h; }

/* Episode start.  The environment's task draws its start uniform from the thread's RandGen first (always consumed:
 * pendulum.cpp:97-103 and siblings), then Agent::start.  Returns the action's value. */
int refag_start(void *h, int test, const double *obs, double *action)
{
  RefAgent *a = (RefAgent *)h;
  try
  {
    (void)RandGen::get();
    Action act;
    Observation o(vec(obs, a->c.obs_dims));
    if (test) a->test_agent.start(o, &act); else a->agent.start(o, &act);
    *action = act.v[0];
    return 0;
  }
  catch (std::exception &e) { fprintf(stderr, "refag_start: %s\n", e.what()); return 1; }
}

/* One step of online_learning.cpp:186-221: terminal == 2 -> Agent::end (absorbing), else Agent::step.  tau = 1
 * (discrete_time = 1, modeled.cpp:209-212). */
int refag_step(void *h, int test, const double *obs, double reward, int terminal, double *action)
{
  RefAgent *a = (RefAgent *)h;
  try
  {
    Agent *ag = test ? (Agent *)&a->test_agent : (Agent *)&a->agent;
    Observation o(vec(obs, a->c.obs_dims), terminal == 2);
    if (terminal == 2) { ag->end(1., o, reward); *action = 0; }
    else { Action act; ag->step(1., o, reward, &act); *action = act.v[0]; }
    return 0;
  }
  catch (std::exception &e) { fprintf(stderr, "refag_step: %s\n", e.what()); return 1; }
}

/* which = 0: Q / critic table, 1: actor table */
int refag_table(void *h, int which, double *out)
{
  RefAgent *a = (RefAgent *)h;
  const LargeVector &p = which ? a->actor_repr.params() : a->repr.params();
  for (size_t i = 0; i < p.size(); ++i) out[i] = p[i];
  return (int)p.size();
}

/* the projector alone: T indices of `in` (dims values) */
int refag_project(void *h, int actor, const double *in, int dims, uint32_t *index)
{
  RefAgent *a = (RefAgent *)h;
  ProjectionPtr p = (actor ? a->actor_proj : a->proj).project(vec(in, dims));
  IndexProjection *ip = dynamic_cast<IndexProjection *>(p.get());
  if (!ip) return -1;
  for (size_t i = 0; i < ip->indices.size(); ++i) index[i] = (uint32_t)ip->indices[i];
  return (int)ip->indices.size();
}

/* Closed loop over the reference's own pieces: task start / actuate / observe / evaluate (environments/*.cpp),
 * DynamicalModel::step (modeled.cpp:254-276) and the agents above.  What is RESTATED here, because ModeledEnvironment and
 * OnlineLearningExperiment read their configuration through the YAML-side configurator tree: the order of calls inside
 * ModeledEnvironment::start / step (modeled.cpp:132-213: actuate, model step, observe, evaluate; discrete_time = 1 -> tau 1)
 * and the episode loop of online_learning.cpp:154-222 (test episode every test_interval + 1 trials; terminal == 2 ->
 * Agent::end, else Agent::step).  Runs n_steps agent-steps (test steps included), as orc_run_total does.
 * out[0] = learning steps, out[1] = trials, out[2] = return of the last finished episode; state_out = the environment state. */
int refag_run(void *h, long n_steps, double *out, double *state_out)
{
  RefAgent *a = (RefAgent *)h;
  const orc_config &c = a->c;
  try
  {
    if (!a->task)
    {
      Configuration cfg;
      cfg.set("timeout", c.timeout); cfg.set("randomization", c.randomization);
      cfg.set("end_stop", 1);
      cfg.set("shaping", 0); cfg.set("gamma", 1.0); cfg.set("action_penalty", 0); cfg.set("end_stop_penalty", c.end_stop_penalty);
      cfg.set("wrap_angle", 1);
      Dynamics *d;
      if (c.dynamics == ORC_DYN_PENDULUM) { a->pd.configure(cfg); a->pt.configure(cfg); d = &a->pd; a->task = &a->pt; }
      else if (c.dynamics == ORC_DYN_CART_POLE) { a->cpd.configure(cfg); a->cpt.configure(cfg); d = &a->cpd; a->task = &a->cpt; }
      else { a->cdpd.configure(cfg); a->cdpt.configure(cfg); d = &a->cdpd; a->task = &a->cdpt; }
      Configuration mc;
      mc.set("dynamics", (void *)static_cast<Configurable *>(d));
      mc.set("control_step", c.control_step);
      mc.set("integration_steps", c.integration_steps);
      a->model.configure(mc);
    }
    double last_return = 0;
    for (long k = 0; k < n_steps; ++k)
    {
      if (!a->in_episode)
      {
        a->test = c.test_interval >= 0 && a->tt % (uint64_t)(c.test_interval + 1) == (uint64_t)c.test_interval;
        int terminal;
        a->task->start(a->test, &a->state);
        Observation obs;
        a->task->observe(a->state, &obs, &terminal);
        if (a->test) a->test_agent.start(obs, &a->action); else a->agent.start(obs, &a->action);
        a->total_reward = 0;
        a->in_episode = true;
      }
      Agent *ag = a->test ? (Agent *)&a->test_agent : (Agent *)&a->agent;
      Vector next, actuation;
      int terminal;
      double reward;
      a->task->actuate(a->state, a->state, a->action, &actuation);
      a->model.step(a->state, actuation, &next);
      Observation obs;
      a->task->observe(next, &obs, &terminal);
      a->task->evaluate(a->state, a->action, next, &reward);
      a->state = next;
      a->total_reward += reward;
      if (terminal == 2) ag->end(1., obs, reward); else ag->step(1., obs, reward, &a->action);
      if (!a->test) a->ss++;
      if (terminal) { a->in_episode = false; a->tt++; last_return = a->total_reward; }
    }
    out[0] = (double)a->ss; out[1] = (double)a->tt; out[2] = last_return;
    for (size_t i = 0; i < a->state.size(); ++i) state_out[i] = a->state[i];
    return 0;
  }
  catch (std::exception &e) { fprintf(stderr, "refag_run: %s\n", e.what()); return 1; }
}

/* The parameters a registered type of the reference requests (Configurable::request, configurable.h:738-742), one line per
 * parameter: name|type|default value|mutability (P rovided, S ystem, C onfiguration, O nline)|optional.  For the host shims'
 * API-surface test (tests/test_host_request_parity.py).  Returns the number of characters written, -1 for an unknown type. */
int refag_requests(const char *type, char *out, int cap)
{
  try
  {
    Configurable *obj = ConfigurableFactory::create(type);
    if (!obj) return -1;
    ConfigurationRequest req;
    obj->request("", &req);
    std::ostringstream oss;
    for (size_t i = 0; i < req.size(); ++i)
      oss << req[i].name << "|" << req[i].type << "|" << req[i].value << "|" << "PSCO"[(int)req[i].mutability] << "|" << (req[i].optional ? 1 : 0) << "\n";
    delete obj;
    const std::string s = oss.str();
    if ((int)s.size() + 1 > cap) return -2;
    memcpy(out, s.c_str(), s.size() + 1);
    return (int)s.size();
  }
  catch (std::exception &e) { fprintf(stderr, "refag_requests: %s\n", e.what()); return -3; }
}

/* DynamicalModel::step (modeled.cpp:254-276: RK4 over the dynamics' eom) of the reference's own modeled.cpp around its own
 * dynamics classes: dyn 0 pendulum (3 states), 1 cart-pole (5), 2 cart-double-pole (7).  Returns tau, or -1. */
double refag_model_step(int dyn, double control_step, int integration_steps, const double *state, double actuation, double *next)
{
  try
  {
    static PendulumDynamics pd; static CartPoleDynamics cpd; static CartDoublePoleDynamics cdpd;
    static bool configured = false;
    if (!configured)
    {
      Configuration cfg;
      cfg.set("end_stop", 1);                         /* dynamics/cart_pole default (cart_pole.h:52) */
      pd.configure(cfg); cpd.configure(cfg); cdpd.configure(cfg);
      configured = true;
    }
    Dynamics *d = dyn == 0 ? (Dynamics *)&pd : dyn == 1 ? (Dynamics *)&cpd : (Dynamics *)&cdpd;
    const int S = dyn == 0 ? 3 : dyn == 1 ? 5 : 7;
    DynamicalModel model;
    Configuration cfg;
    cfg.set("dynamics", (void *)static_cast<Configurable *>(d));
    cfg.set("control_step", control_step);
    cfg.set("integration_steps", integration_steps);
    model.configure(cfg);
    Vector out;
    const double tau = model.step(vec(state, S), VectorConstructor(actuation), &out);
    for (size_t i = 0; i < out.size(); ++i) next[i] = out[i];
    return tau;
  }
  catch (std::exception &e) { fprintf(stderr, "refag_model_step: %s\n", e.what()); return -1; }
}

}  /* extern "C" */
