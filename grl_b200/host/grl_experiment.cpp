// grl_experiment.cpp -- experiment/online_learning with an `agents` count: the reference's episode loop
// (online_learning.cpp:154-279) executed for `agents` independent agents per launch of the fused kernel.
#include "grl_objects.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <numeric>

namespace grl {

REGISTER_CONFIGURABLE(OnlineLearningExperiment)
REGISTER_CONFIGURABLE(MultiExperiment)

void OnlineLearningExperiment::request(ConfigurationRequest *config)
{
  config->push_back(CRP("runs", "Number of separate learning runs to perform", runs_, CRP::Configuration, 1));
  config->push_back(CRP("run_offset", "Run offset to start at", run_offset_));
  config->push_back(CRP("trials", "Number of episodes per learning run", trials_));
  config->push_back(CRP("steps", "Number of steps per learning run", steps_));
  config->push_back(CRP("rate", "Control step frequency in Hz", rate_, CRP::Online));
  config->push_back(CRP("test_interval", "Number of episodes in between test trials", test_interval_, CRP::Configuration, -1));
  config->push_back(CRP("test_trials", "Number of test trials per interval", test_trials_, CRP::Configuration, 1));
  config->push_back(CRP("output", "Output base filename", output_));
  config->push_back(CRP("environment", "environment", "Environment in which the agent acts", environment_));
  config->push_back(CRP("agent", "agent", "Agent", agent_));
  config->push_back(CRP("test_agent", "agent", "Agent to use in test trials", agent_, true));
  config->push_back(CRP("load_file", "Load policy filename", load_file_));
  config->push_back(CRP("save_every", "Save policy to 'output' at the end of event", save_every_));
  // new on the batched path
  config->push_back(CRP("agents", "Independent agents learning side by side on the GPU (per run)", agents_, CRP::Configuration, 1));
  config->push_back(CRP("checkpoint_agents", "Agents whose tables load_file reads and save_every writes (0: all; agent k's file is '<name>-agent<k>.dat', agent 0's has the reference's own name)", checkpoint_agents_, CRP::Configuration, 0));
  config->push_back(CRP("steps_per_launch", "Agent-steps per agent between two reports", steps_per_launch_, CRP::Configuration, 1));
  // Several GPUs of one box (what MultiExperiment's instances on threads are to the reference, experiments/multi.cpp:44-75):
  // the agents are sharded over `devices` in contiguous ranges of global agent ids, one context per device, and the
  // per-agent results are all-gathered over NCCL before every report.  Results are keyed by global agent id, so the
  // output does not depend on the number of devices.
  config->push_back(CRP("device", "CUDA device ordinal (-1: every visible device whose count divides `agents`)", device_, CRP::Configuration, -1));
  config->push_back(CRP("devices", "CUDA device ordinals to shard the agents over (overrides `device`)", Vector(), CRP::Configuration));
  // Hyper-parameter grid on the agents axis: what MultiExperiment `instances` (experiments/multi.cpp:38-60) and a
  // grlo `parameters` list (bin/optimize.yaml) do with one process per setting.  For the agent with global id g:
  // lambda is constant over blocks of 32 consecutive agents (a warp sweeps as long as its longest trace), block
  // b = g / 32 runs lambda[b mod |lambda|]; alpha and epsilon cycle through the agents that share a lambda
  // (k = (b / |lambda|) * 32 + g mod 32: alpha[k mod |alpha|], epsilon[(k / |alpha|) mod |epsilon|]).  Agents with the
  // same setting are its repetitions; cell id = alpha index + |alpha| (lambda index + |lambda| epsilon index).
  config->push_back(CRP("sweep_alpha", "Learning rates of the grid (empty: the predictor's alpha)", Vector(), CRP::Configuration));
  config->push_back(CRP("sweep_lambda", "Trace decay rates of the grid (empty: the predictor's lambda)", Vector(), CRP::Configuration));
  config->push_back(CRP("sweep_epsilon", "Exploration rates of the grid (empty: the sampler's epsilon)", Vector(), CRP::Configuration));
}

void OnlineLearningExperiment::configure(Configuration &config)
{
  agent_ = dynamic_cast<Agent *>(config["agent"].ptr());
  test_agent_ = dynamic_cast<Agent *>(config["test_agent"].ptr());
  environment_ = dynamic_cast<Environment *>(config["environment"].ptr());
  runs_ = config["runs"]; run_offset_ = config["run_offset"]; trials_ = config["trials"]; steps_ = config["steps"];
  rate_ = config["rate"]; test_interval_ = config["test_interval"]; test_trials_ = config["test_trials"];
  output_ = config["output"].str(); load_file_ = config["load_file"].str(); save_every_ = config["save_every"].str();
  agents_ = config["agents"]; steps_per_launch_ = config["steps_per_launch"]; device_ = config["device"];
  devices_ = config["devices"].v();
  checkpoint_agents_ = config["checkpoint_agents"];
  if (checkpoint_agents_ < 0) throw bad_param("experiment/online_learning:checkpoint_agents");
  for (double d : devices_) if (d < 0 || d != floor(d)) throw bad_param("experiment/online_learning:devices");
  if (!devices_.empty() && agents_ % (int)devices_.size()) throw bad_param("experiment/online_learning:{agents,devices}");
  sweep_alpha_ = config["sweep_alpha"].v(); sweep_lambda_ = config["sweep_lambda"].v(); sweep_epsilon_ = config["sweep_epsilon"].v();
  for (double v : sweep_alpha_) if (!(v > 0)) throw bad_param("experiment/online_learning:sweep_alpha");
  for (double v : sweep_lambda_) if (!(v >= 0 && v <= 1)) throw bad_param("experiment/online_learning:sweep_lambda");
  for (double v : sweep_epsilon_) if (!(v >= 0 && v <= 1)) throw bad_param("experiment/online_learning:sweep_epsilon");
  if (runs_ < 1) throw bad_param("experiment/online_learning:runs");
  if (agents_ < 1) throw bad_param("experiment/online_learning:agents");
  if (steps_per_launch_ < 1) throw bad_param("experiment/online_learning:steps_per_launch");
  if (test_trials_ != 1) throw bad_param("experiment/online_learning:test_trials");
  // online_learning.cpp:62: {"never", "run", "test", "trial"}.  On the batched path a report (every steps_per_launch
  // agent-steps) is the unit of progress: "trial" saves after every report, "test" after every report once the agents
  // have finished a new test episode, into `output`-run<r>-trial<min finished episodes>-...
  if (save_every_ != "never" && save_every_ != "run" && save_every_ != "test" && save_every_ != "trial")
    throw bad_param("experiment/online_learning:save_every");
  if (test_interval_ >= 0 && !test_agent_) throw bad_param("experiment/online_learning:test_agent");
}

void OnlineLearningExperiment::reconfigure(const Configuration &config)
{
  config.get("rate", rate_);
  if (config.has("identity")) identity = config["identity"].str();
}

// Collect the parameters of the CUDA objects into one grlcu_config (no device needed).
void OnlineLearningExperiment::assemble()
{
  if (assembled_) return;
  CudaBatchedEnvironment *env = dynamic_cast<CudaBatchedEnvironment *>(environment_);
  TDAgent *td = dynamic_cast<TDAgent *>(agent_);
  if (!env) throw Exception(path() + ": environment must be an environment/model/cuda_batched");
  if (!td || !td->policy)
    throw Exception(path() + ": agent must be an agent/td with a mapping/policy/discrete/value/q or mapping/policy/action policy");
  Policy *pol = td->policy;
  CudaTileCodingProjector *proj = dynamic_cast<CudaTileCodingProjector *>(pol->projector);
  CudaLinearRepresentation *rep = dynamic_cast<CudaLinearRepresentation *>(pol->representation);
  if (!proj) throw Exception(path() + ": policy projector must be a projector/tile_coding/cuda");
  if (!rep) throw Exception(path() + ": policy representation must be a representation/linear/cuda");
  grlcu_config &c = path_.cfg;

  if (CudaSARSAPredictor *pred = dynamic_cast<CudaSARSAPredictor *>(td->predictor))
  {
    // SARSA(lambda): QPolicy and predictor share one (obs, action) projector and one Q table
    QPolicy *qp = dynamic_cast<QPolicy *>(pol);
    if (!qp) throw Exception(path() + ": predictor/sarsa/cuda needs a mapping/policy/discrete/value/q policy");
    if (pred->projector() != pol->projector || pred->representation() != pol->representation)
      throw Exception(path() + ": predictor and policy must share one projector and one representation");
    if (!qp->discretizer || !qp->sampler) throw Exception(path() + ": policy needs a discretizer and a sampler");
    if (CudaQPredictor *q = dynamic_cast<CudaQPredictor *>(pred))
      if (q->discretizer() != qp->discretizer) throw Exception(path() + ": predictor/q/cuda must maximise over the policy's discretizer");
    if (CudaExpectedSARSAPredictor *es = dynamic_cast<CudaExpectedSARSAPredictor *>(pred))
    {
      QPolicy *tp = es->targetPolicy();
      if (tp->projector != pol->projector || tp->representation != pol->representation || tp->discretizer != qp->discretizer)
        throw Exception(path() + ": the expected-SARSA target policy must evaluate the learning agent's discretizer, projector and representation");
      if (!sweep_epsilon_.empty() && tp->sampler == qp->sampler)
        throw Exception(path() + ": sweep_epsilon with an on-policy expected-SARSA target is not supported (one target epsilon per context)");
    }
    if (test_agent_ && test_agent_->policy)
    {
      QPolicy *tp = dynamic_cast<QPolicy *>(test_agent_->policy);
      if (!tp || tp->projector != pol->projector || tp->representation != pol->representation || tp->discretizer != qp->discretizer)
        throw Exception(path() + ": test_agent must evaluate the learning agent's discretizer, projector and representation");
      if (!dynamic_cast<GreedySampler *>(tp->sampler)) throw Exception(path() + ": test_agent sampler must be sampler/greedy");
    }
    env->contribute(&path_);
    proj->contribute(&path_);
    rep->contribute(&path_);
    pred->contribute(&path_);
    c.n_actions = (int)qp->discretizer->values.size();
    for (int i = 0; i < c.n_actions; ++i) c.actions[i] = qp->discretizer->values[i];
    c.epsilon = qp->sampler->epsilon;
    if (dynamic_cast<SoftmaxSampler *>(qp->sampler))
    {
      if (!sweep_epsilon_.empty()) throw bad_param("experiment/online_learning:sweep_epsilon");   // no epsilon to sweep
      c.sampler = GRLCU_SAMPLER_SOFTMAX;
      c.softmax_tau = qp->sampler->tau;
    }
    env->attach(&path_); proj->attach(&path_); rep->attach(&path_);
    checkpoint_reps_ = {rep};
  }
  else if (CudaACPredictor *ac = dynamic_cast<CudaACPredictor *>(td->predictor))
  {
    // Actor-critic (cfg/cart_double_pole/ac.yaml): ActionPolicy + predictor/ac/action{critic: predictor/critic/td}.
    // The actor's and the critic's projectors must be configured identically: the tile hash has no per-object
    // seed, so one index set addresses both per-agent tables (grlcu_ac.cuh).
    ActionPolicy *ap = dynamic_cast<ActionPolicy *>(pol);
    if (!ap) throw Exception(path() + ": predictor/ac/action/cuda needs a mapping/policy/action policy");
    if (ac->projector() != pol->projector || ac->representation() != pol->representation)
      throw Exception(path() + ": predictor/ac/action/cuda must update the policy's projector and representation");
    CudaTileCodingProjector *cproj = dynamic_cast<CudaTileCodingProjector *>(ac->critic()->projector());
    CudaLinearRepresentation *crep = dynamic_cast<CudaLinearRepresentation *>(ac->critic()->representation());
    if (crep == rep) throw Exception(path() + ": actor and critic need separate representations");
    if (cproj->tilings() != proj->tilings() || cproj->memory() != proj->memory() ||
        cproj->resolution() != proj->resolution() || cproj->wrapping() != proj->wrapping())
      throw Exception(path() + ": the critic's projector must have the actor projector's tilings, memory, resolution and wrapping");
    if (rep->memory() != proj->memory() || crep->memory() != proj->memory())
      throw bad_param("representation/linear/cuda:memory (or matching projector)");
    if (rep->outputMin() != ap->min || rep->outputMax() != ap->max)
      throw Exception(path() + ": the action representation's output_min/max must be the policy's (the task's action limits)");
    if (rep->initMin() != crep->initMin() || rep->initMax() != crep->initMax())
      throw Exception(path() + ": actor and critic representations must share init_min/init_max");
    if (test_agent_ && test_agent_->policy)
    {
      ActionPolicy *tp = dynamic_cast<ActionPolicy *>(test_agent_->policy);
      if (!tp || tp->projector != pol->projector || tp->representation != pol->representation)
        throw Exception(path() + ": test_agent must evaluate the learning agent's projector and representation");
      if (tp->sigma != 0.) throw Exception(path() + ": test_agent policy must not explore (sigma: [ ])");
    }
    env->contribute(&path_);
    cproj->contribute(&path_);
    crep->contribute(&path_);
    ac->contribute(&path_);
    c.n_actions = 0;
    c.epsilon = 0;
    c.sigma = ap->sigma;
    c.action_min = ap->min[0]; c.action_max = ap->max[0];
    for (size_t i = 0; i < proj->resolution().size(); ++i)
    {
      c.actor_resolution[i] = proj->resolution()[i];
      c.actor_wrapping[i] = proj->wrapping()[i];
    }
    env->attach(&path_); proj->attach(&path_); cproj->attach(&path_);
    rep->attach(&path_, true); crep->attach(&path_);
    checkpoint_reps_ = {rep, crep};
  }
  else
    throw Exception(path() + ": agent predictor must be a predictor/sarsa/cuda or a predictor/ac/action/cuda");
  // the trace ring is sized for the largest lambda of the grid (grlcu_set_hyper checks it)
  for (double v : sweep_lambda_) c.lambda = std::max(c.lambda, v);
  if (!sweep_epsilon_.empty() && c.algo == GRLCU_ALGO_AC) throw bad_param("experiment/online_learning:sweep_epsilon");
  c.test_interval = test_interval_;
  c.seed = seed;
  path_.agents = agents_;
  path_.device = device_ < 0 ? 0 : device_;
  path_.devices.clear();
  for (double d : devices_) path_.devices.push_back((int)d);
  assembled_ = true;
}

std::vector<double> OnlineLearningExperiment::run()
{
  assemble();
  if (!trials_ && !steps_)
    throw Exception(path() + ": set `trials` or `steps` (the batched experiment does not run until interrupted)");
  std::vector<double> curve;
  const size_t n = (size_t)agents_;
  std::vector<double> last_return(n), last_test(n);
  std::vector<uint64_t> ss(n);
  std::vector<uint32_t> episodes(n);
  grlcu_results res;
  memset(&res, 0, sizeof(res));
  res.last_return = last_return.data(); res.last_test_return = last_test.data();
  res.steps = ss.data(); res.episodes = episodes.data();

  for (int rr = run_offset_; rr < runs_ + run_offset_; ++rr)
  {
    // every run is a fresh set of agents with its own Philox streams (global agent ids rr*agents ...)
    path_.release();
    path_.first_agent_id = (long long)rr * agents_;
    if (devices_.empty() && device_ < 0)
    {
      // default: every visible device, as many as divide the agents evenly
      int G = std::max(1, grlcu_device_count());
      while (G > 1 && agents_ % G) --G;
      path_.devices.clear();
      for (int g = 0; g < G; ++g) path_.devices.push_back(g);
    }
    path_.context();
    const int per = path_.perContext();
    const size_t na = std::max<size_t>(sweep_alpha_.size(), 1), nl = std::max<size_t>(sweep_lambda_.size(), 1),
                 ne = std::max<size_t>(sweep_epsilon_.size(), 1), cells = na * nl * ne;
    const bool sweeping = !sweep_alpha_.empty() || !sweep_lambda_.empty() || !sweep_epsilon_.empty();
    std::vector<size_t> cell_of(n, 0);
    if (sweeping)
    {
      std::vector<double> al(n), la(n), ep(n);
      for (size_t a = 0; a < n; ++a)
      {
        const unsigned long long g = (unsigned long long)path_.first_agent_id + a;
        const unsigned long long blk = g / 32, k = (blk / nl) * 32 + g % 32;
        const size_t ia = (size_t)(k % na), il = (size_t)(blk % nl), ie = (size_t)((k / na) % ne);
        cell_of[a] = ia + na * (il + nl * ie);
        al[a] = sweep_alpha_.empty() ? path_.cfg.alpha : sweep_alpha_[ia];
        la[a] = sweep_lambda_.empty() ? path_.cfg.lambda : sweep_lambda_[il];
        ep[a] = sweep_epsilon_.empty() ? path_.cfg.epsilon : sweep_epsilon_[ie];
      }
      for (size_t g = 0; g < path_.ctxs.size(); ++g)
        CudaPath::check(grlcu_set_hyper(path_.ctxs[g], al.data() + g * per, la.data() + g * per, ep.data() + g * per));
    }
    std::ofstream ofs, cfs;
    if (!output_.empty())
    {
      std::ostringstream oss;
      oss << output_ << "-" << rr << identity << ".txt";
      ofs.open(oss.str().c_str());
      if (sweeping)
      {
        std::ostringstream css;
        css << output_ << "-" << rr << identity << "-cells.txt";
        cfs.open(css.str().c_str());
        cfs << "# report cell alpha lambda epsilon agents mean p10 p50 p90   (of the last "
            << (test_interval_ >= 0 ? "test" : "learning") << " return of the cell's agents)" << std::endl;
      }
    }
    size_t report = 0;
    uint32_t saved_tests = 0;
    if (!load_file_.empty())
    {
      Configuration loadconfig;
      loadconfig.set("action", "load");
      // online_learning.cpp:141-150: `load_file` + "-", with "$run" replaced by the run number
      std::string load_file = load_file_ + "-";
      for (size_t at = load_file.find("$run"); at != std::string::npos; at = load_file.find("$run", at))
        load_file.replace(at, 4, std::to_string(rr));
      loadconfig.set("file", load_file);
      loadconfig.set("agents", checkpoint_agents_ ? std::min(agents_, checkpoint_agents_) : agents_);
      for (Representation *r : checkpoint_reps_) r->reconfigure(loadconfig);
    }
    auto start = std::chrono::steady_clock::now();
    {
      std::ostringstream h;
      h << std::setw(15) << "Trial" << std::setw(15) << "Steps" << std::setw(15) << "Episode reward" << std::setw(15)
        << "Reward min" << std::setw(15) << "Reward max" << std::setw(15) << "Wall time";
      INFO(h.str());
    }
    for (;;)
    {
      CudaPath::check(grlcu_group_run(path_.group, steps_per_launch_));
      double collective_ms = 0;
      CudaPath::check(grlcu_group_results(path_.group, &res, &collective_ms));   // the only collective: NCCL all-gather
      collective_ms_total_ += collective_ms;
      const uint64_t min_ss = *std::min_element(ss.begin(), ss.end());
      const uint32_t min_tt = *std::min_element(episodes.begin(), episodes.end());
      const std::vector<double> &r = test_interval_ >= 0 ? last_test : last_return;
      const double mean = std::accumulate(r.begin(), r.end(), 0.) / n;
      const double mean_tt = std::accumulate(episodes.begin(), episodes.end(), 0.) / n;
      const double mean_ss = std::accumulate(ss.begin(), ss.end(), 0.) / n;
      const double wall = std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - start).count() / 1e6;
      std::ostringstream oss;
      oss << std::setw(15) << std::fixed << std::setprecision(1) << mean_tt << std::setw(15) << std::setprecision(0) << mean_ss
          << std::setw(15) << std::setprecision(3) << mean << std::setw(15) << *std::min_element(r.begin(), r.end())
          << std::setw(15) << *std::max_element(r.begin(), r.end()) << std::setw(15) << wall;
      INFO(oss.str());
      if (ofs.is_open()) ofs << oss.str() << std::endl;
      curve.push_back(mean);
      if (cfs.is_open())
      {
        // learning curve per grid cell: mean and percentiles over the cell's repetitions
        std::vector<std::vector<double>> by_cell(cells);
        for (size_t a = 0; a < n; ++a) by_cell[cell_of[a]].push_back(r[a]);
        for (size_t cell = 0; cell < cells; ++cell)
        {
          std::vector<double> &v = by_cell[cell];
          if (v.empty()) continue;
          std::sort(v.begin(), v.end());
          auto pct = [&](double q) { return v[std::min(v.size() - 1, (size_t)(q * v.size()))]; };
          cfs << report << " " << cell << " " << (sweep_alpha_.empty() ? path_.cfg.alpha : sweep_alpha_[cell % na]) << " "
              << (sweep_lambda_.empty() ? path_.cfg.lambda : sweep_lambda_[(cell / na) % nl]) << " "
              << (sweep_epsilon_.empty() ? path_.cfg.epsilon : sweep_epsilon_[cell / (na * nl)]) << " " << v.size() << " "
              << std::setprecision(9) << std::accumulate(v.begin(), v.end(), 0.) / v.size() << " " << pct(0.1) << " " << pct(0.5) << " "
              << pct(0.9) << std::endl;
        }
      }
      // online_learning.cpp:281-291: save the policy every trial / every test trial
      const bool new_test = test_interval_ >= 0 && min_tt / (uint32_t)(test_interval_ + 1) > saved_tests;
      if ((save_every_ == "trial" || (save_every_ == "test" && new_test)) && !output_.empty())
      {
        saved_tests = min_tt / (uint32_t)(test_interval_ >= 0 ? test_interval_ + 1 : 1);
        std::ostringstream soss;
        soss << output_ << "-run" << rr << "-trial" << min_tt << "-";
        Configuration saveconfig;
        saveconfig.set("action", "save");
        saveconfig.set("file", soss.str());
        saveconfig.set("agents", checkpoint_agents_ ? std::min(agents_, checkpoint_agents_) : agents_);
        for (Representation *r : checkpoint_reps_) r->reconfigure(saveconfig);
      }
      ++report;
      if ((steps_ && min_ss >= (uint64_t)steps_) || (trials_ && min_tt >= (uint32_t)trials_)) break;
    }
    if (save_every_ == "run" && !output_.empty())
    {
      std::ostringstream oss;
      oss << output_ << "-run" << rr << "-";
      Configuration saveconfig;
      saveconfig.set("action", "save");
      saveconfig.set("file", oss.str());
      saveconfig.set("agents", checkpoint_agents_ ? std::min(agents_, checkpoint_agents_) : agents_);
      for (Representation *r : checkpoint_reps_) r->reconfigure(saveconfig);
    }
    if (ofs.is_open())
    {
      // per-agent summary: the batched analogue of one learning curve per `grld` process
      ofs << "# agent steps episodes last_return last_test_return" << std::endl;
      for (size_t a = 0; a < n; ++a)
        ofs << "# " << a << " " << ss[a] << " " << episodes[a] << " " << std::setprecision(17) << last_return[a] << " " << last_test[a] << std::endl;
    }
  }
  return curve;
}

}  // namespace grl

namespace grl {

void MultiExperiment::request(ConfigurationRequest *config)
{
  config->push_back(CRP("instances", "Number of experiments to run in parallel", instances_, CRP::Configuration, 1));
  config->push_back(CRP("experiment", "experiment", "Experiment to run", prototype_));
}

void MultiExperiment::configure(Configuration &config)
{
  instances_ = config["instances"];
  prototype_ = dynamic_cast<OnlineLearningExperiment *>(config["experiment"].ptr());
  if (instances_ < 1) throw bad_param("experiment/multi:instances");
  if (!prototype_) throw Exception(path() + ": experiment must be an experiment/online_learning");
  // Every clone must own its tables (see the class comment): the policy has to be configured inside the prototype.
  Agent *agent = prototype_->agent();
  const std::string inside = prototype_->path() + "/";
  if (!agent || !agent->policy || agent->policy->path().compare(0, inside.size(), inside) != 0)
    throw Exception(path() + ": the instances share one policy (" + (agent && agent->policy ? agent->policy->path() : std::string("?")) +
                    " lies outside " + prototype_->path() + "): the reference's clones would then update ONE table from several "
                    "threads without synchronisation (cfg/pendulum/multi_sarsa_tc.yaml), which is not what independent agents "
                    "compute.  Configure the policy inside the experiment, or use `agents` / `sweep_*` of experiment/online_learning");
  prototype_->setAgents(instances_ * prototype_->agents());
}

std::vector<double> MultiExperiment::run()
{
  return prototype_->run();
}

}  // namespace grl
