// grl_objects.cpp -- parameter carriers and the four CUDA types (see grl_objects.h).
#include "grl_objects.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace grl {

// ---- carriers ---------------------------------------------------------------------------------------
REGISTER_CONFIGURABLE(PendulumDynamics)
REGISTER_CONFIGURABLE(CartPoleDynamics)
REGISTER_CONFIGURABLE(CartDoublePoleDynamics)
REGISTER_CONFIGURABLE(DynamicalModel)
REGISTER_CONFIGURABLE(PendulumSwingupTask)
REGISTER_CONFIGURABLE(CartPoleSwingupTask)
REGISTER_CONFIGURABLE(CartDoublePoleBalancingTask)
REGISTER_CONFIGURABLE(UniformDiscretizer)
REGISTER_CONFIGURABLE(GreedySampler)
REGISTER_CONFIGURABLE(EpsilonGreedySampler)
REGISTER_CONFIGURABLE(SoftmaxSampler)
REGISTER_CONFIGURABLE(ReplacingEnumeratedTrace)
REGISTER_CONFIGURABLE(AccumulatingEnumeratedTrace)
REGISTER_CONFIGURABLE(QPolicy)
REGISTER_CONFIGURABLE(ActionPolicy)
REGISTER_CONFIGURABLE(TDAgent)
REGISTER_CONFIGURABLE(FixedAgent)

void CartPoleDynamics::request(ConfigurationRequest *config)
{
  config->push_back(CRP("end_stop", "Simulate end stops (adds position and velocity to state)", 1, CRP::Configuration, 0, 1));
}
void CartPoleDynamics::configure(Configuration &config)
{
  // cart_pole.cpp:37-40; the device code implements the default (end_stop = 1, cart_pole.h:52)
  if ((int)config["end_stop"] != 1) throw bad_param("dynamics/cart_pole:end_stop");
}

void DynamicalModel::request(ConfigurationRequest *config)
{
  config->push_back(CRP("control_step", "double.control_step", "Control step time", tau, CRP::Configuration, 0.00001, DBL_MAX));
  config->push_back(CRP("integration_steps", "Number of integration steps per control step", steps, CRP::Configuration, 1));
  config->push_back(CRP("dynamics", "dynamics", "Equations of motion", dynamics));
}
void DynamicalModel::configure(Configuration &config)
{
  dynamics = (Dynamics *)config["dynamics"].ptr();
  tau = config["control_step"];
  steps = config["integration_steps"];
  if (!(tau >= 0.00001)) throw bad_param("model/dynamical:control_step");
  if (steps < 1) throw bad_param("model/dynamical:integration_steps");
}

static void provide_limits(Configuration &config, Task &t, int obs_dims, double reward_min, double reward_max)
{
  config.set("observation_dims", obs_dims);
  config.set("observation_min", t.observation_min);
  config.set("observation_max", t.observation_max);
  config.set("action_dims", 1);
  config.set("action_min", t.action_min);
  config.set("action_max", t.action_max);
  config.set("reward_min", reward_min);
  config.set("reward_max", reward_max);
  t.reward_min = reward_min; t.reward_max = reward_max;
}

static void request_limits(ConfigurationRequest *config)
{
  // Task::request (environment.h): what a task provides to the rest of the tree
  config->push_back(CRP("observation_dims", "int.observation_dims", "Number of observation dimensions", CRP::Provided));
  config->push_back(CRP("observation_min", "vector.observation_min", "Lower limit on observations", CRP::Provided));
  config->push_back(CRP("observation_max", "vector.observation_max", "Upper limit on observations", CRP::Provided));
  config->push_back(CRP("action_dims", "int.action_dims", "Number of action dimensions", CRP::Provided));
  config->push_back(CRP("action_min", "vector.action_min", "Lower limit on actions", CRP::Provided));
  config->push_back(CRP("action_max", "vector.action_max", "Upper limit on actions", CRP::Provided));
  config->push_back(CRP("reward_min", "double.reward_min", "Lower limit on immediate reward", CRP::Provided));
  config->push_back(CRP("reward_max", "double.reward_max", "Upper limit on immediate reward", CRP::Provided));
}

void PendulumSwingupTask::request(ConfigurationRequest *config)
{
  request_limits(config);
  config->push_back(CRP("timeout", "Episode timeout", 2.99, CRP::Configuration, 0., DBL_MAX));
  config->push_back(CRP("randomization", "Level of start state randomization", 0., CRP::Configuration, 0., 1.));
}
void PendulumSwingupTask::configure(Configuration &config)
{
  // pendulum.cpp:78-91
  dynamics_id = GRLCU_DYN_PENDULUM;
  timeout = config["timeout"];
  randomization = config["randomization"];
  if (timeout < 0) throw bad_param("task/pendulum/swingup:timeout");
  if (randomization < 0 || randomization > 1) throw bad_param("task/pendulum/swingup:randomization");
  observation_min = {0., -12 * M_PI}; observation_max = {2 * M_PI, 12 * M_PI};
  action_min = {-3}; action_max = {3};
  provide_limits(config, *this, 2, -5 * pow(M_PI, 2) - 0.1 * pow(12 * M_PI, 2) - 1 * pow(3, 2), 0);
}

void CartPoleSwingupTask::request(ConfigurationRequest *config)
{
  request_limits(config);
  config->push_back(CRP("timeout", "Episode timeout", 9.99, CRP::Configuration, 0., DBL_MAX));
  config->push_back(CRP("randomization", "Start state randomization", 0, CRP::Online, 0, 1));     // an int in the reference (cart_pole.cpp:114)
  config->push_back(CRP("shaping", "Whether to use reward shaping", 0, CRP::Configuration, 0, 1));
  config->push_back(CRP("gamma", "Discount rate for reward shaping", 1., CRP::Configuration, 0., 1.));
  config->push_back(CRP("end_stop_penalty", "Terminate episode with penalty when end stop is reached", 1, CRP::Configuration, 0, 1));
  config->push_back(CRP("action_penalty", "Penalize applied torque", 0, CRP::Configuration, 0, 1));
}
void CartPoleSwingupTask::configure(Configuration &config)
{
  // cart_pole.cpp:122-147; the device code implements shaping = 0, action_penalty = 0 (the defaults of
  // cart_pole.h:75) and both values of end_stop_penalty (cfg/cart_pole/ac_tc.yaml switches it off)
  dynamics_id = GRLCU_DYN_CART_POLE;
  timeout = config["timeout"];
  randomization = config["randomization"];
  if ((int)config["shaping"] != 0) throw bad_param("task/cart_pole/swingup:shaping");
  end_stop_penalty = config["end_stop_penalty"];
  if ((int)config["action_penalty"] != 0) throw bad_param("task/cart_pole/swingup:action_penalty");
  observation_min = {-2.4, 0., -10.0, -5 * M_PI}; observation_max = {2.4, 2 * M_PI, 10.0, 5 * M_PI};
  action_min = {-15.}; action_max = {15.};
  provide_limits(config, *this, 4, -2 * pow(2.4, 2) - 0.1 * pow(10, 2) - pow(M_PI, 2) - 0.1 * pow(5 * M_PI, 2) - end_stop_penalty * 10000, 0);
}

void CartDoublePoleBalancingTask::request(ConfigurationRequest *config)
{
  request_limits(config);
  config->push_back(CRP("timeout", "Episode timeout", 9.99, CRP::Configuration, 0., DBL_MAX));     // the reference's default (cart_double_pole.cpp:236); its cfg files set 4.99
}
void CartDoublePoleBalancingTask::configure(Configuration &config)
{
  // cart_double_pole.cpp:225-246; start() always draws theta1 = 0.01 U - 0.005 (no randomization parameter)
  dynamics_id = GRLCU_DYN_CART_DOUBLE_POLE;
  timeout = config["timeout"];
  randomization = 0;
  if (timeout < 0) throw bad_param("task/cart_double_pole/balancing:timeout");
  observation_min = {-2.4, -0.7, -0.7, -5.0, -5.0, -5.0}; observation_max = {2.4, 0.7, 0.7, 5.0, 5.0, 5.0};
  action_min = {-20.}; action_max = {20.};
  provide_limits(config, *this, 6, 0, 6.2);
}

void UniformDiscretizer::request(const std::string &role, ConfigurationRequest *config)
{
  const std::string suffix = role == "action" ? "action" : role == "observation" ? "observation" : "";
  config->push_back(CRP("min", suffix.empty() ? "vector" : "vector." + suffix + "_min", "Lower limit", Vector(), CRP::System));
  config->push_back(CRP("max", suffix.empty() ? "vector" : "vector." + suffix + "_max", "Upper limit", Vector(), CRP::System));
  config->push_back(CRP("steps", "Discretization steps per dimension", Vector(), CRP::Configuration));
}
void UniformDiscretizer::configure(Configuration &config)
{
  // uniform.cpp:60-94: value k of dimension d = min + k*(max-min)/(steps-1)
  min = config["min"].v(); max = config["max"].v(); steps = config["steps"].v();
  if (min.size() != max.size() || min.empty()) throw bad_param("discretizer/uniform:{min,max}");
  if (steps.size() != min.size()) throw bad_param("discretizer/uniform:steps");
  if (min.size() != 1) throw Exception(path() + ": the batched CUDA path supports one action dimension");
  if (steps[0] < 1 || steps[0] > GRLCU_MAX_ACTIONS) throw bad_param("discretizer/uniform:steps");
  values.clear();
  const int n = (int)steps[0];
  for (int k = 0; k < n; ++k)
    values.push_back(n > 1 ? min[0] + k * (max[0] - min[0]) / (n - 1) : min[0]);
}

void GreedySampler::request(ConfigurationRequest *config)
{
  config->push_back(CRP("rand_max", "In case of multiple maxima, pick at random", 0, CRP::Configuration, 0, 1));
}
void EpsilonGreedySampler::request(ConfigurationRequest *config)
{
  config->push_back(CRP("rand_max", "In case of multiple maxima, pick at random", 0, CRP::Configuration, 0, 1));
  config->push_back(CRP("epsilon", "Exploration rate (can be defined per action)", Vector{0.05}, CRP::Online));
  config->push_back(CRP("decay_rate", "Multiplicative decay factor per episode", 1., CRP::Configuration));
  config->push_back(CRP("decay_min", "Minimum decay (eps_min = eps*decay_min)", 0., CRP::Configuration));
}
void EpsilonGreedySampler::configure(Configuration &config)
{
  Vector e = config["epsilon"].v();
  if (e.size() != 1) throw Exception(path() + ": the batched CUDA path supports a single epsilon");
  if ((double)config["decay_rate"] != 1.) throw bad_param("sampler/epsilon_greedy:decay_rate");
  epsilon = e[0];
}

void SoftmaxSampler::request(ConfigurationRequest *config)
{
  // softmax.cpp:32-35
  config->push_back(CRP("tau", "Temperature of Boltzmann distribution", 1., CRP::Online, 0.000001, 100.));
}
void SoftmaxSampler::configure(Configuration &config)
{
  tau = config["tau"];
  if (!(tau >= 0.000001 && tau <= 100.)) throw bad_param("sampler/softmax:tau");
}

void QPolicy::request(ConfigurationRequest *config)
{
  config->push_back(CRP("discretizer", "discretizer.action", "Action discretizer", discretizer));
  config->push_back(CRP("projector", "projector.pair", "Projects observation-action pairs onto representation space", projector));
  config->push_back(CRP("representation", "representation.value/action", "Action-value representation", representation));
  config->push_back(CRP("sampler", "sampler", "Samples actions from action-values", sampler));
}
void QPolicy::configure(Configuration &config)
{
  discretizer = (UniformDiscretizer *)config["discretizer"].ptr();
  projector = (Projector *)config["projector"].ptr();
  representation = (Representation *)config["representation"].ptr();
  sampler = (Sampler *)config["sampler"].ptr();
}

void ActionPolicy::request(ConfigurationRequest *config)
{
  // policies/action.cpp:39-55
  config->push_back(CRP("sigma", "Standard deviation of Gaussian exploration distribution", Vector(), CRP::Configuration));
  config->push_back(CRP("theta", "Ornstein-Uhlenbeck friction term (1=pure Gaussian noise)", Vector(), CRP::Configuration));
  config->push_back(CRP("decay_rate", "Multiplicative decay factor per episode", 1., CRP::Configuration));
  config->push_back(CRP("decay_min", "Minimum decay (sigma_min = sigma*decay_min)", 0., CRP::Configuration));
  config->push_back(CRP("renormalize", "Renormalize representation output from [-1, 1] to [min, max]", 0, CRP::Configuration, 0, 1));
  config->push_back(CRP("output_min", "vector.action_min", "Lower limit on outputs", Vector(), CRP::System));
  config->push_back(CRP("output_max", "vector.action_max", "Upper limit on outputs", Vector(), CRP::System));
  config->push_back(CRP("projector", "projector.observation", "Projects observations onto representation space", projector));
  config->push_back(CRP("representation", "representation.action", "Action representation", representation));
}
void ActionPolicy::configure(Configuration &config)
{
  // policies/action.cpp:57-97
  projector = (Projector *)config["projector"].ptr();
  representation = (Representation *)config["representation"].ptr();
  min = config["output_min"].v(); max = config["output_max"].v();
  if (min.size() != max.size() || min.empty()) throw bad_param("policy/action:{output_min,output_max}");
  if (min.size() != 1) throw Exception(path() + ": the batched CUDA path supports one action dimension");
  Vector s = config["sigma"].v(), th = config["theta"].v();
  if (s.empty()) s = Vector{0.};
  if (s.size() != 1) throw bad_param("policy/action:sigma");
  if (th.empty()) th = Vector{1.};
  if (th.size() != 1) throw bad_param("policy/action:theta");
  if (th[0] != 1.) throw bad_param("policy/action:theta");                 // Ornstein-Uhlenbeck noise: not on the device
  if ((double)config["decay_rate"] != 1.) throw bad_param("policy/action:decay_rate");
  if ((int)config["renormalize"] != 0) throw bad_param("policy/action:renormalize");
  sigma = s[0];
}

void TDAgent::request(ConfigurationRequest *config)
{
  config->push_back(CRP("policy", "mapping/policy", "Control policy", policy));
  config->push_back(CRP("predictor", "predictor", "Value function predictor", predictor));
}
void TDAgent::configure(Configuration &config)
{
  policy = dynamic_cast<Policy *>(config["policy"].ptr());
  predictor = (Predictor *)config["predictor"].ptr();
}
void FixedAgent::request(ConfigurationRequest *config)
{
  config->push_back(CRP("policy", "mapping/policy", "Control policy", policy));
}
void FixedAgent::configure(Configuration &config) { policy = dynamic_cast<Policy *>(config["policy"].ptr()); }

// ---- shared device context ------------------------------------------------------------------------------
CudaPath::CudaPath() { memset(&cfg, 0, sizeof(cfg)); cfg.test_interval = -1; cfg.seed = 1; cfg.end_stop_penalty = 1; }
CudaPath::~CudaPath() { release(); }

void CudaPath::check(int rc)
{
  if (rc != GRLCU_OK) throw Exception(std::string("libgrlcu: ") + grlcu_last_error());
}

grlcu_ctx *CudaPath::context()
{
  if (ctx) return ctx;
  if (!have_env || !have_projector || !have_representation || !have_predictor)
    throw Exception("CUDA path used before environment/model/cuda_batched, projector/tile_coding/cuda, "
                    "representation/linear/cuda and predictor/sarsa/cuda were all configured in one experiment");
  // Agents shard trivially over the devices: device g owns the global ids [first + g*agents/G, first + (g+1)*agents/G)
  std::vector<int> devs = devices.empty() ? std::vector<int>{device} : devices;
  const int G = (int)devs.size();
  if (agents % G) throw Exception("`agents` must be a multiple of the number of devices");
  const int per = agents / G;
  try
  {
    for (int g = 0; g < G; ++g)
    {
      grlcu_ctx *c = nullptr;
      check(grlcu_create(&cfg, per, first_agent_id + (long long)g * per, devs[g], nullptr, &c));
      ctxs.push_back(c);
    }
    check(grlcu_group_create(ctxs.data(), G, &group));
  }
  catch (...)
  {
    release();
    throw;
  }
  ctx = ctxs[0];
  return ctx;
}

grlcu_ctx *CudaPath::contextOf(int agent, int *local)
{
  context();
  const int per = agents / (int)ctxs.size();
  if (agent < 0 || agent >= agents) throw Exception("agent index out of range");
  *local = agent % per;
  return ctxs[agent / per];
}

void CudaPath::release()
{
  if (group) grlcu_group_destroy(group);
  group = nullptr;
  for (grlcu_ctx *c : ctxs) grlcu_destroy(c);
  ctxs.clear();
  ctx = nullptr;
}

}  // namespace grl
