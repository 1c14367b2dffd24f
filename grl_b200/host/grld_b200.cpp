// grld_b200 -- deployer for the batched CUDA path, command-line compatible with the reference's grld
// (base/src/deployer.cpp:35-150): grld_b200 [-v] [-s seed] [-c] <yaml file>.
//   -s seed   Philox seed of every agent stream (the reference seeds srand48 with it)
//   -v        more log output (repeatable)
//   -c        check only: load, instantiate and print the resolved C-ABI configuration as JSON; no GPU needed
#include "grl_objects.h"

#include <cstdlib>
#include <cstring>
#include <unistd.h>

using namespace grl;

static void print_config_json(const grlcu_config &c, int agents)
{
  auto vec = [](const double *v, int n) { std::string s = "["; char b[64]; for (int i = 0; i < n; ++i) { snprintf(b, sizeof(b), "%s%.17g", i ? ", " : "", v[i]); s += b; } return s + "]"; };
  const int dims = c.obs_dims + 1;
  printf("{\"agents\": %d, \"dynamics\": %d, \"integration_steps\": %d, \"control_step\": %.17g, \"timeout\": %.17g, "
         "\"randomization\": %.17g, \"tilings\": %d, \"memory\": %d, \"obs_dims\": %d, \"n_actions\": %d, "
         "\"resolution\": %s, \"wrapping\": %s, \"actions\": %s, \"alpha\": %.17g, \"gamma\": %.17g, \"lambda\": %.17g, "
         "\"epsilon\": %.17g, \"trace\": %d, \"test_interval\": %d, \"init_min\": %.17g, \"init_max\": %.17g, \"seed\": %llu, \"algo\": %d, "
         "\"sigma\": %.17g, \"actor_alpha\": %.17g, \"action_min\": %.17g, \"action_max\": %.17g, \"actor_resolution\": %s, \"actor_wrapping\": %s, \"target_epsilon\": %.17g, \"end_stop_penalty\": %d, \"sampler\": %d, \"softmax_tau\": %.17g, \"safe\": %d}\n",
         agents, c.dynamics, c.integration_steps, c.control_step, c.timeout, c.randomization, c.tilings, c.memory, c.obs_dims,
         c.n_actions, vec(c.resolution, dims).c_str(), vec(c.wrapping, dims).c_str(), vec(c.actions, c.n_actions).c_str(),
         c.alpha, c.gamma, c.lambda, c.epsilon, c.trace, c.test_interval, c.init_min, c.init_max, (unsigned long long)c.seed, c.algo,
         c.sigma, c.actor_alpha, c.action_min, c.action_max, vec(c.actor_resolution, c.algo == GRLCU_ALGO_AC ? c.obs_dims : 0).c_str(),
         vec(c.actor_wrapping, c.algo == GRLCU_ALGO_AC ? c.obs_dims : 0).c_str(), c.target_epsilon, c.end_stop_penalty, c.sampler, c.softmax_tau, c.safe);
}

// -i: drive the one-agent role interfaces of the CUDA objects (Environment::start/step, Projector::project,
// Representation::read/write/update: the signatures of environment.h:48-54, projector.h:52-69,
// representation.h:51-92) the way a GRL agent would, and print what they returned as one JSON line.
static void print_vec(const char *name, const Vector &v, bool last = false)
{
  printf("\"%s\": [", name);
  for (size_t i = 0; i < v.size(); ++i) printf("%s%.17g", i ? ", " : "", v[i]);
  printf("]%s", last ? "" : ", ");
}

static int exercise_interfaces(Configurator &configurator, OnlineLearningExperiment *experiment)
{
  experiment->resolved();                                     // attaches the objects to the experiment's device context
  Environment *env = dynamic_cast<Environment *>(configurator.object("experiment/environment"));
  Projector *proj = dynamic_cast<Projector *>(configurator.object("experiment/agent/policy/projector"));
  Representation *rep = dynamic_cast<Representation *>(configurator.object("experiment/agent/policy/representation"));
  if (!env || !proj || !rep) { ERROR("-i needs experiment/environment and experiment/agent/policy/{projector,representation}"); return 1; }
  const grlcu_config &cfg = experiment->resolved();
  Observation obs, obs2;
  env->start(0, &obs);
  Vector in = obs;
  if (cfg.algo != GRLCU_ALGO_AC) in.push_back(cfg.actions[cfg.n_actions - 1]);   // extend(obs, action) (q.cpp:94-107)
  ProjectionPtr p = proj->project(in);
  Vector idx(p->indices.begin(), p->indices.end());
  Vector q0, q1, q2;
  rep->read(p, &q0);
  rep->write(p, Vector{q0[0] + 1.}, Vector{0.5});             // linear.cpp:186-196: every weight += 0.5 * 1
  rep->read(p, &q1);
  rep->update(p, Vector{-0.25});
  rep->read(p, &q2);
  double reward = 0; int terminal = 0;
  const double tau = env->step(Vector{cfg.algo == GRLCU_ALGO_AC ? 1.5 : cfg.actions[cfg.n_actions - 1]}, &obs2, &reward, &terminal);
  printf("{");
  print_vec("obs", obs); print_vec("index", idx); print_vec("q", Vector{q0[0], q1[0], q2[0]}); print_vec("next_obs", obs2);
  printf("\"reward\": %.17g, \"terminal\": %d, \"tau\": %.17g}\n", reward, terminal, tau);
  return 0;
}

int main(int argc, char **argv)
{
  unsigned long long seed = 1;
  bool check = false, interfaces = false;
  int c;
  while ((c = getopt(argc, argv, "vcis:")) != -1)
    switch (c)
    {
      case 'v': grl_log_verbosity__++; break;
      case 'c': check = true; break;
      case 'i': interfaces = true; break;
      case 's': seed = strtoull(optarg, nullptr, 10); break;
      default: return 1;
    }
  if (optind > argc - 1)
  {
    ERROR("Usage: " << std::endl << "  " << argv[0] << " [-v] [-c] [-i] [-s seed] <yaml file>");
    return 1;
  }
  if ((check || interfaces) && grl_log_verbosity__ == 3) grl_log_verbosity__ = 1;     // keep stdout to the JSON line
  try
  {
    NOTICE("Loading configuration from '" << argv[optind] << "'");
    std::shared_ptr<YamlNode> doc = loadYAMLFile(argv[optind]);
    Configurator configurator;
    NOTICE("Instantiating configuration");
    configurator.instantiate(*doc);
    Experiment *top = dynamic_cast<Experiment *>(configurator.object("experiment"));
    OnlineLearningExperiment *experiment = dynamic_cast<OnlineLearningExperiment *>(top);
    if (MultiExperiment *multi = dynamic_cast<MultiExperiment *>(top)) experiment = multi->prototype();   // instances -> agents
    if (!top || !experiment)
    {
      ERROR("YAML configuration does not specify an experiment");
      return 1;
    }
    experiment->seed = seed;
    if (check)
    {
      print_config_json(experiment->resolved(), experiment->agents());
      return 0;
    }
    if (interfaces) return exercise_interfaces(configurator, experiment);
    NOTICE("Starting experiment");
    top->run();
    if (experiment->collectiveMs() > 0) NOTICE("NCCL all-gather of the per-agent results: " << experiment->collectiveMs() << " ms in total");
    NOTICE("Exiting");
  }
  catch (Exception &e)
  {
    ERROR(e.what());
    return 1;
  }
  return 0;
}
