// grl_cuda_objects.cpp -- environment/model/cuda_batched, projector/tile_coding/cuda,
// representation/linear/cuda, predictor/sarsa/cuda: Configurable shims over the C ABI of include/grlcu.h.
#include "grl_objects.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace grl {

REGISTER_CONFIGURABLE(CudaBatchedEnvironment)
REGISTER_CONFIGURABLE(CudaTileCodingProjector)
REGISTER_CONFIGURABLE(CudaLinearRepresentation)
REGISTER_CONFIGURABLE(CudaSARSAPredictor)
REGISTER_CONFIGURABLE(CudaQPredictor)
REGISTER_CONFIGURABLE(CudaExpectedSARSAPredictor)
REGISTER_CONFIGURABLE(CudaTDPredictor)
REGISTER_CONFIGURABLE(CudaACPredictor)

// ---- environment/model/cuda_batched (stands in for environment/modeled, modeled.cpp:35-121) ---------------
void CudaBatchedEnvironment::request(ConfigurationRequest *config)
{
  config->push_back(CRP("discrete_time", "Always report unit step time", discrete_time_, CRP::Configuration, 0, 1));
  config->push_back(CRP("window", "Number of observations to concatenate", 1, CRP::Configuration, 1, 10));
  config->push_back(CRP("stride", "Time steps between concatenated observations", 1, CRP::Configuration, 1, 100));
  config->push_back(CRP("delta", "Action delta for differential actions", Vector(), CRP::Configuration));
  config->push_back(CRP("model", "model", "Environment model", model_));
  config->push_back(CRP("task", "task", "Task to perform in the environment (should match model)", task_));
  config->push_back(CRP("observation_dims", "int.observation_dims", "Number of observation dimensions", CRP::Provided));
  config->push_back(CRP("observation_min", "vector.observation_min", "Lower limit on observations", CRP::Provided));
  config->push_back(CRP("observation_max", "vector.observation_max", "Upper limit on observations", CRP::Provided));
  config->push_back(CRP("action_dims", "int.action_dims", "Number of action dimensions", CRP::Provided));
  config->push_back(CRP("action_min", "vector.action_min", "Lower limit on actions", CRP::Provided));
  config->push_back(CRP("action_max", "vector.action_max", "Upper limit on actions", CRP::Provided));
  config->push_back(CRP("reward_min", "vector.reward_min", "Lower limit on immediate reward", CRP::Provided));
  config->push_back(CRP("reward_max", "vector.reward_max", "Upper limit on immediate reward", CRP::Provided));
}

void CudaBatchedEnvironment::configure(Configuration &config)
{
  discrete_time_ = config["discrete_time"];
  model_ = dynamic_cast<DynamicalModel *>(config["model"].ptr());
  task_ = dynamic_cast<Task *>(config["task"].ptr());
  if (!model_) throw Exception(path() + ": model must be a model/dynamical");
  if (!task_) throw Exception(path() + ": task must be one of the batched tasks (pendulum/swingup, cart_pole/swingup, cart_double_pole/balancing)");
  if ((int)config["window"] != 1) throw bad_param("environment/model/cuda_batched:window");
  if ((int)config["stride"] != 1) throw bad_param("environment/model/cuda_batched:stride");
  if (!config["delta"].v().empty()) throw bad_param("environment/model/cuda_batched:delta");
  // the fused kernel reports tau = 1 to the agent (discrete_time_ = 1, modeled.cpp:209-212)
  if (discrete_time_ != 1) throw bad_param("environment/model/cuda_batched:discrete_time");
  if (!model_->dynamics || model_->dynamics->id != task_->dynamics_id)
    throw Exception(path() + ": task does not match the model's dynamics");
  config.set("observation_dims", (int)task_->observation_min.size());
  config.set("observation_min", task_->observation_min);
  config.set("observation_max", task_->observation_max);
  config.set("action_dims", 1);
  config.set("action_min", task_->action_min);
  config.set("action_max", task_->action_max);
  config.set("reward_min", task_->reward_min);          // modeled.cpp:112-113
  config.set("reward_max", task_->reward_max);
}

void CudaBatchedEnvironment::reconfigure(const Configuration &) { }

void CudaBatchedEnvironment::contribute(CudaPath *p) const
{
  p->cfg.dynamics = task_->dynamics_id;
  p->cfg.integration_steps = model_->steps;
  p->cfg.control_step = model_->tau;
  p->cfg.timeout = task_->timeout;
  p->cfg.randomization = task_->randomization;
  p->cfg.end_stop_penalty = task_->end_stop_penalty;
  p->cfg.obs_dims = (int)task_->observation_min.size();
  p->have_env = true;
}

// Environment::start / step for one agent (environment.h:48-54): the same device code as the fused loop.
void CudaBatchedEnvironment::start(int test, Observation *obs)
{
  if (!path_) throw Exception(path() + ": not attached to an experiment");
  const int S = path_->cfg.dynamics == GRLCU_DYN_PENDULUM ? 3 : path_->cfg.dynamics == GRLCU_DYN_CART_POLE ? 5 : 7;
  const int OBS = path_->cfg.obs_dims;
  const double u = 0.5;                     // agent-0 view: centre draw; the batched path draws from Philox
  state_.assign(S, 0.);
  obs->assign(OBS, 0.);
  CudaPath::check(grlcu_task_start(path_->context(), 1, test, &u, state_.data(), obs->data(), nullptr));
}

double CudaBatchedEnvironment::step(const Action &action, Observation *obs, double *reward, int *terminal)
{
  if (!path_ || state_.empty()) throw Exception(path() + ": step() before start()");
  if (action.size() != 1) throw Exception(path() + ": expected a 1-dimensional action");
  Vector next(state_.size());
  obs->assign(path_->cfg.obs_dims, 0.);
  int32_t term = 0;
  CudaPath::check(grlcu_model_step(path_->context(), 1, state_.data(), action.data(), next.data(), obs->data(), reward, &term));
  state_ = next;
  *terminal = term;
  return 1.;                               // discrete_time_
}

// ---- projector/tile_coding/cuda (tile_coding.cpp:35-80) -------------------------------------------------
void CudaTileCodingProjector::request(ConfigurationRequest *config)
{
  config->push_back(CRP("tilings", "Number of tilings", tilings_));
  config->push_back(CRP("memory", "int.memory", "Hash table size (entries per agent)", memory_));
  config->push_back(CRP("safe", "Collision detection (0=off, 1=claim on write, 2=claim always)", safe_, CRP::Configuration, 0, 2));
  config->push_back(CRP("resolution", "Size of a single tile", resolution_));
  config->push_back(CRP("wrapping", "vector.wrapping", "Wrapping boundaries (must be multiple of resolution)", wrapping_));
}

void CudaTileCodingProjector::configure(Configuration &config)
{
  tilings_ = config["tilings"];
  memory_ = config["memory"];
  safe_ = config["safe"];
  resolution_ = config["resolution"].v();
  wrapping_ = config["wrapping"].v();
  if (tilings_ != 16) throw bad_param("projector/tile_coding/cuda:tilings");       // kernels instantiated for 16
  if (memory_ < 1) throw bad_param("projector/tile_coding/cuda:memory");
  if (safe_ < 0 || safe_ > 2) throw bad_param("projector/tile_coding/cuda:safe");   // 1, 2: per-agent collision-claim table (SARSA only)
  if (resolution_.empty() || resolution_.size() > GRLCU_MAX_DIMS) throw bad_param("projector/tile_coding/cuda:resolution");
  if (wrapping_.empty()) wrapping_ = Vector(resolution_.size(), 0.);
  if (wrapping_.size() != resolution_.size()) throw bad_param("projector/tile_coding:wrapping");
  for (size_t ii = 0; ii < resolution_.size(); ++ii)
  {
    const double w = wrapping_[ii] * (tilings_ / resolution_[ii]);
    if (fabs(w - round(w)) > 0.001)
    {
      ERROR("Scaled wrapping for dimension " << ii << " (" << w << ") is not an integer");
      throw bad_param("projector/tile_coding:wrapping");
    }
  }
}

void CudaTileCodingProjector::contribute(CudaPath *p) const
{
  p->cfg.tilings = tilings_;
  p->cfg.memory = memory_;
  p->cfg.safe = safe_;
  for (size_t i = 0; i < resolution_.size(); ++i) { p->cfg.resolution[i] = resolution_[i]; p->cfg.wrapping[i] = wrapping_[i]; }
  p->have_projector = true;
}

ProjectionPtr CudaTileCodingProjector::project(const Vector &in) const
{
  if (!path_) throw Exception(path() + ": not attached to an experiment");
  if (in.size() != resolution_.size()) throw bad_param("projector/tile_coding:resolution");
  std::vector<uint32_t> index(tilings_);
  CudaPath::check(grlcu_project(path_->context(), 1, in.data(), nullptr, index.data()));
  ProjectionPtr p(new IndexProjection());
  p->indices.assign(index.begin(), index.end());
  return p;
}

// ---- representation/linear/cuda (linear.cpp:35-125, representation.h:201-263) ----------------------------
void CudaLinearRepresentation::request(const std::string &, ConfigurationRequest *config)
{
  config->push_back(CRP("init_min", "Lower initial value limit", Vector{0.}));
  config->push_back(CRP("init_max", "Upper initial value limit", Vector{1.}));
  config->push_back(CRP("memory", "int.memory", "Feature vector size", memory_, CRP::System));
  config->push_back(CRP("outputs", "Number of outputs", outputs_, CRP::System));
  config->push_back(CRP("output_min", "Lower output limit", Vector(), CRP::System));
  config->push_back(CRP("output_max", "Upper output limit", Vector(), CRP::System));
  config->push_back(CRP("limit", "Limit parameters to same range as outputs", limit_, CRP::Configuration, 0, 1));
  config->push_back(CRP("interval", "Target representation update interval", 0, CRP::Configuration));
  config->push_back(CRP("tau", "Target representation update strength", 1., CRP::Configuration));     // a double in the reference (representation.h:181)
  config->push_back(CRP("file_suffix", "Checkpoint file suffix (default: configuration path)", "", CRP::Configuration));
}

void CudaLinearRepresentation::configure(Configuration &config)
{
  init_min_ = config["init_min"].v(); init_max_ = config["init_max"].v();
  output_min_ = config["output_min"].v(); output_max_ = config["output_max"].v();
  memory_ = config["memory"]; outputs_ = config["outputs"]; limit_ = config["limit"];
  if (memory_ < 1) throw bad_param("representation/linear/cuda:memory");
  if (outputs_ != 1) throw bad_param("representation/linear/cuda:outputs");
  if (init_min_.size() != 1 || init_max_.size() != 1) throw bad_param("representation/linear/cuda:{init_min,init_max}");
  // unset limits mean +-DBL_MAX (linear.cpp:88-98): the value tables.  Set limits are what the actor-critic
  // policy's action representation carries (= the task's action_min/max); the experiment checks the role.
  if (output_min_.size() != output_max_.size() || output_min_.size() > 1) throw bad_param("representation/linear/cuda:{output_min,output_max}");
  if (!output_min_.empty() && limit_ != 1) throw bad_param("representation/linear/cuda:limit");
  if ((int)config["interval"] != 0) throw bad_param("representation/linear/cuda:interval");
}

void CudaLinearRepresentation::contribute(CudaPath *p) const
{
  if (p->have_projector && p->cfg.memory != memory_)
    throw bad_param("representation/linear/cuda:memory (or matching projector)");
  if (!output_min_.empty()) throw bad_param("representation/linear/cuda:{output_min,output_max}");   // value table: unlimited
  p->cfg.init_min = init_min_[0]; p->cfg.init_max = init_max_[0];
  p->have_representation = true;
}

double CudaLinearRepresentation::read(const ProjectionPtr &projection, Vector *result) const
{
  if (!path_) throw Exception(path() + ": not attached to an experiment");
  std::vector<uint64_t> idx(projection->indices.begin(), projection->indices.end());
  double v = 0;
  CudaPath::check((actor_ ? grlcu_read_actor : grlcu_read)(path_->context(), 0, (int)idx.size(), idx.data(), &v));
  result->assign(1, v);
  return v;
}

void CudaLinearRepresentation::update(const ProjectionPtr projection, const Vector &delta)
{
  if (!path_) throw Exception(path() + ": not attached to an experiment");
  if (delta.size() != 1) throw Exception(path() + ": expected a 1-dimensional delta");
  std::vector<uint64_t> idx(projection->indices.begin(), projection->indices.end());
  CudaPath::check((actor_ ? grlcu_update_actor : grlcu_update)(path_->context(), 0, (int)idx.size(), idx.data(), delta[0]));
}

void CudaLinearRepresentation::write(const ProjectionPtr projection, const Vector &target, const Vector &alpha)
{
  // linear.cpp:186-196: delta = alpha*(target - read(projection)); update(projection, delta)
  Vector value;
  read(projection, &value);
  update(projection, Vector{alpha[0] * (target[0] - value[0])});
}

void CudaLinearRepresentation::reconfigure(const Configuration &config)
{
  if (!config.has("action") || !path_) return;
  const std::string action = config["action"].str();
  if (action == "reset")
  {
    if (path_->ctx && !actor_)
      for (grlcu_ctx *c : path_->ctxs) CudaPath::check(grlcu_reset(c, GRLCU_RESET_TABLES));   // resets both tables of an AC context
    return;
  }
  if (action != "save" && action != "load") return;
  // representation.h:201-263: raw little-endian doubles, `file` + path with '/' -> '_' + ".dat".  Agent 0 uses
  // exactly that name (so reference grld can load it); agent k > 0 appends "-agent<k>" before ".dat".
  std::string base = config.has("file") ? config["file"].str() : std::string();
  // The reference's object paths start at the root, "/experiment/agent/policy/representation", so its file names carry a
  // leading '_' ("<file>_experiment_agent_policy_representation.dat"; seen in the files the reference's own deployer
  // writes, oracle/_ref/grld_ref with save_every: run).  This host's paths have no leading slash.
  std::string cfg_path = path();
  if (cfg_path.empty() || cfg_path[0] != '/') cfg_path = "/" + cfg_path;
  std::replace(cfg_path.begin(), cfg_path.end(), '/', '_');
  const int count = std::min(path_->agents, config.has("agents") ? (int)config["agents"] : 1);
  std::vector<double> p(memory_);
  for (int a = 0; a < count; ++a)
  {
    const std::string file = base + cfg_path + (a ? "-agent" + std::to_string(a) : std::string()) + ".dat";
    int local = 0;
    grlcu_ctx *ctx = path_->contextOf(a, &local);
    if (action == "save")
    {
      CudaPath::check((actor_ ? grlcu_get_actor_table : grlcu_get_table)(ctx, local, p.data()));
      FILE *f = fopen(file.c_str(), "wb");
      if (!f) { WARNING("Could not open '" << file << "' for writing"); return; }
      fwrite(p.data(), sizeof(double), p.size(), f);
      fclose(f);
    }
    else
    {
      FILE *f = fopen(file.c_str(), "rb");
      if (!f) { WARNING("Could not open '" << file << "' for reading"); return; }
      fseek(f, 0, SEEK_END);
      if (ftell(f) != (long)(p.size() * sizeof(double))) { WARNING("Configuration mismatch for '" << file << "'"); fclose(f); return; }
      fseek(f, 0, SEEK_SET);
      if (fread(p.data(), sizeof(double), p.size(), f) != p.size()) { WARNING("Could not read '" << file << "'"); fclose(f); return; }
      fclose(f);
      CudaPath::check((actor_ ? grlcu_set_actor_table : grlcu_set_table)(ctx, local, p.data()));
    }
  }
}

// ---- predictor/sarsa/cuda (sarsa.cpp:35-67) -------------------------------------------------------------
void CudaSARSAPredictor::request(ConfigurationRequest *config)
{
  config->push_back(CRP("alpha", "Learning rate", alpha_));
  config->push_back(CRP("gamma", "Discount rate", gamma_));
  config->push_back(CRP("lambda", "Trace decay rate", lambda_));
  config->push_back(CRP("projector", "projector.pair", "Projects observation-action pairs onto representation space", projector_));
  config->push_back(CRP("representation", "representation.value/action", "Q-value representation", representation_));
  config->push_back(CRP("trace", "trace", "Trace of projections", trace_, true));
}

void CudaSARSAPredictor::configure(Configuration &config)
{
  projector_ = (Projector *)config["projector"].ptr();
  representation_ = (Representation *)config["representation"].ptr();
  trace_ = dynamic_cast<Trace *>(config["trace"].ptr());
  alpha_ = config["alpha"]; gamma_ = config["gamma"]; lambda_ = config["lambda"];
  if (!dynamic_cast<CudaTileCodingProjector *>(projector_)) throw Exception(path() + ": projector must be a projector/tile_coding/cuda");
  if (!dynamic_cast<CudaLinearRepresentation *>(representation_)) throw Exception(path() + ": representation must be a representation/linear/cuda");
  if (config["trace"].ptr() && !trace_) throw Exception(path() + ": trace must be trace/enumerated/replacing or trace/enumerated/accumulating");
}

void CudaSARSAPredictor::reconfigure(const Configuration &) { }

void CudaSARSAPredictor::contribute(CudaPath *p) const
{
  p->cfg.alpha = alpha_; p->cfg.gamma = gamma_; p->cfg.lambda = lambda_;
  p->cfg.trace = trace_ ? trace_->kind : GRLCU_TRACE_NONE;
  p->cfg.algo = GRLCU_ALGO_SARSA;
  p->have_predictor = true;
}

// ---- predictor/q/cuda (advantage.cpp:35-69), predictor/expected_sarsa/cuda (sarsa.cpp:134-165) ------------------
void CudaQPredictor::request(ConfigurationRequest *config)
{
  CudaSARSAPredictor::request(config);
  config->push_back(CRP("discretizer", "discretizer.action", "Action discretizer", discretizer_));
}

void CudaQPredictor::configure(Configuration &config)
{
  CudaSARSAPredictor::configure(config);
  discretizer_ = dynamic_cast<UniformDiscretizer *>(config["discretizer"].ptr());
  if (!discretizer_) throw Exception(path() + ": discretizer must be a discretizer/uniform");
}

void CudaQPredictor::contribute(CudaPath *p) const
{
  CudaSARSAPredictor::contribute(p);
  p->cfg.algo = GRLCU_ALGO_Q;
}

void CudaExpectedSARSAPredictor::request(ConfigurationRequest *config)
{
  CudaSARSAPredictor::request(config);
  config->push_back(CRP("policy", "mapping/policy/discrete/value", "Value based target policy", policy_));
}

void CudaExpectedSARSAPredictor::configure(Configuration &config)
{
  CudaSARSAPredictor::configure(config);
  policy_ = dynamic_cast<QPolicy *>(config["policy"].ptr());
  if (!policy_ || !policy_->sampler) throw Exception(path() + ": policy must be a mapping/policy/discrete/value/q with a sampler");
}

void CudaExpectedSARSAPredictor::contribute(CudaPath *p) const
{
  CudaSARSAPredictor::contribute(p);
  p->cfg.algo = GRLCU_ALGO_EXPECTED_SARSA;
  // GreedySampler::distribution (greedy.cpp:88-98) or EpsilonGreedySampler::distribution (greedy.cpp:220-238)
  if (dynamic_cast<SoftmaxSampler *>(policy_->sampler))
    throw Exception(path() + ": the expected-SARSA target policy must use sampler/greedy or sampler/epsilon_greedy");
  p->cfg.target_epsilon = dynamic_cast<EpsilonGreedySampler *>(policy_->sampler) ? policy_->sampler->epsilon : -1.;
}

// ---- predictor/critic/td/cuda (predictors/td.cpp:35-66) ----------------------------------------------------
void CudaTDPredictor::request(ConfigurationRequest *config)
{
  config->push_back(CRP("alpha", "Learning rate", alpha_));
  config->push_back(CRP("gamma", "Discount rate", gamma_));
  config->push_back(CRP("lambda", "Trace decay rate", lambda_));
  config->push_back(CRP("projector", "projector.observation", "Projects observations onto representation space", projector_));
  config->push_back(CRP("representation", "representation.value/state", "State value representation", representation_));
  config->push_back(CRP("trace", "trace", "Trace of projections", trace_, true));
}

void CudaTDPredictor::configure(Configuration &config)
{
  projector_ = (Projector *)config["projector"].ptr();
  representation_ = (Representation *)config["representation"].ptr();
  trace_ = dynamic_cast<Trace *>(config["trace"].ptr());
  alpha_ = config["alpha"]; gamma_ = config["gamma"]; lambda_ = config["lambda"];
  if (!dynamic_cast<CudaTileCodingProjector *>(projector_)) throw Exception(path() + ": projector must be a projector/tile_coding/cuda");
  if (!dynamic_cast<CudaLinearRepresentation *>(representation_)) throw Exception(path() + ": representation must be a representation/linear/cuda");
  if (config["trace"].ptr() && !trace_) throw Exception(path() + ": trace must be trace/enumerated/replacing or trace/enumerated/accumulating");
}

void CudaTDPredictor::contribute(CudaPath *p) const
{
  p->cfg.alpha = alpha_; p->cfg.gamma = gamma_; p->cfg.lambda = lambda_;
  p->cfg.trace = trace_ ? trace_->kind : GRLCU_TRACE_NONE;
}

// ---- predictor/ac/action/cuda (predictors/ac.cpp:35-70) ------------------------------------------------------
void CudaACPredictor::request(ConfigurationRequest *config)
{
  config->push_back(CRP("alpha", "Actor learning rate", alpha_));
  config->push_back(CRP("update_method", "Actor update method", "proportional", CRP::Configuration));
  config->push_back(CRP("step_limit", "Actor exploration step limit", Vector()));
  config->push_back(CRP("projector", "projector.observation", "Projects observations onto actor representation space", projector_));
  config->push_back(CRP("representation", "representation.action", "Action representation", representation_));
  config->push_back(CRP("critic", "predictor/critic", "Critic predictor", critic_));
}

void CudaACPredictor::configure(Configuration &config)
{
  projector_ = (Projector *)config["projector"].ptr();
  representation_ = (Representation *)config["representation"].ptr();
  critic_ = dynamic_cast<CudaTDPredictor *>(config["critic"].ptr());
  alpha_ = config["alpha"];
  if (!critic_) throw Exception(path() + ": critic must be a predictor/critic/td/cuda");
  if (!dynamic_cast<CudaTileCodingProjector *>(projector_)) throw Exception(path() + ": projector must be a projector/tile_coding/cuda");
  if (!dynamic_cast<CudaLinearRepresentation *>(representation_)) throw Exception(path() + ": representation must be a representation/linear/cuda");
  // ac.cpp:84-107: "proportional" is the update on the device; "cacla" and the step limit are not
  if (config["update_method"].str() != "proportional") throw bad_param("predictor/ac/action/cuda:update_method");
  if (!config["step_limit"].v().empty()) throw bad_param("predictor/ac/action/cuda:step_limit");
}

void CudaACPredictor::contribute(CudaPath *p) const
{
  critic_->contribute(p);
  p->cfg.actor_alpha = alpha_;
  p->cfg.algo = GRLCU_ALGO_AC;
  p->have_predictor = true;
}

}  // namespace grl
