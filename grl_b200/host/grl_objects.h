// grl_objects.h -- the GRL objects on the hot path, as Configurable types.
//
// Four new types carry the CUDA path (BASELINE north_star): `environment/model/cuda_batched`,
// `projector/tile_coding/cuda`, `representation/linear/cuda`, `predictor/sarsa/cuda`.  They accept the
// same parameters as the reference types they stand in for (modeled.cpp:35-57, tile_coding.cpp:35-43,
// linear.cpp:35-64, sarsa.cpp:35-46) and keep their role interfaces (one-agent view).  The objects that
// only carry parameters on this path (model, dynamics, task, agent, policy, discretizer, sampler, trace)
// are registered under the reference's own type names so the shipped YAML files load with nothing but the
// four `type:` lines changed.  `experiment/online_learning` gains an `agents` count.
#pragma once
#include "grl_host.h"
#include "../../include/grlcu.h"

namespace grl {

// ---- role interfaces (signatures of the reference's L2 interfaces, one agent per call) -------------------
typedef Vector Observation;
typedef Vector Action;

/// Sparse binary feature vector (reference: IndexProjection, projection.h:68-113).
struct IndexProjection
{
  std::vector<size_t> indices;
  static size_t invalid_index() { return (size_t)-1; }
};
typedef std::shared_ptr<IndexProjection> ProjectionPtr;

class Environment : public Configurable
{
  public:
    virtual void start(int test, Observation *obs) = 0;
    virtual double step(const Action &action, Observation *obs, double *reward, int *terminal) = 0;
};

class Projector : public Configurable
{
  public:
    virtual ProjectionPtr project(const Vector &in) const = 0;
};

class Representation : public Configurable
{
  public:
    virtual double read(const ProjectionPtr &projection, Vector *result) const = 0;
    virtual void write(const ProjectionPtr projection, const Vector &target, const Vector &alpha) = 0;
    virtual void update(const ProjectionPtr projection, const Vector &delta) = 0;
};

class Predictor : public Configurable { };
class Experiment : public Configurable
{
  public:
    virtual std::vector<double> run() = 0;
};

// ---- parameter carriers registered under the reference's type names ---------------------------------------
struct Dynamics : Configurable { int id = -1; };
struct PendulumDynamics : Dynamics { TYPEINFO("dynamics/pendulum", "Pendulum dynamics based on the DCSC MOPS") PendulumDynamics() { id = GRLCU_DYN_PENDULUM; } };
struct CartPoleDynamics : Dynamics
{
  TYPEINFO("dynamics/cart_pole", "Cart-pole dynamics from Barto et al.")
  CartPoleDynamics() { id = GRLCU_DYN_CART_POLE; }
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};
struct CartDoublePoleDynamics : Dynamics { TYPEINFO("dynamics/cart_double_pole", "Cart-double-pole dynamics from Zhong and Rock") CartDoublePoleDynamics() { id = GRLCU_DYN_CART_DOUBLE_POLE; } };

struct DynamicalModel : Configurable
{
  TYPEINFO("model/dynamical", "State transition model that integrates equations of motion")
  double tau = 0.05; int steps = 5; Dynamics *dynamics = nullptr;
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};

struct Task : Configurable
{
  int dynamics_id = -1; double timeout = 0, randomization = 0;
  int end_stop_penalty = 1;      // task/cart_pole/swingup only (cart_pole.h:75)
  Vector observation_min, observation_max, action_min, action_max;
  double reward_min = 0, reward_max = 0;      // re-exported by the environment (modeled.cpp:112-113)
};
struct PendulumSwingupTask : Task
{
  TYPEINFO("task/pendulum/swingup", "Pendulum swing-up task")
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};
struct CartPoleSwingupTask : Task
{
  TYPEINFO("task/cart_pole/swingup", "Cart-pole swing-up task")
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};
struct CartDoublePoleBalancingTask : Task
{
  TYPEINFO("task/cart_double_pole/balancing", "Cart-double-pole balancing task")
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};

struct UniformDiscretizer : Configurable
{
  TYPEINFO("discretizer/uniform", "Uniform discretizer")
  Vector min, max, steps;
  std::vector<double> values;          // one action dimension on this path
  void request(const std::string &role, ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};

struct Sampler : Configurable { double epsilon = 0; double tau = 0; };   // tau > 0: sampler/softmax
struct GreedySampler : Sampler
{
  TYPEINFO("sampler/greedy", "Maximum search")
  void request(ConfigurationRequest *config) override;
};
struct EpsilonGreedySampler : Sampler
{
  TYPEINFO("sampler/epsilon_greedy", "Maximum search with an epsilon chance of randomness")
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};

struct SoftmaxSampler : Sampler
{
  TYPEINFO("sampler/softmax", "Softmax (Boltzmann) sampler")
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};

struct Trace : Configurable { int kind = GRLCU_TRACE_NONE; };
struct ReplacingEnumeratedTrace : Trace { TYPEINFO("trace/enumerated/replacing", "Replacing eligibility trace using a queue of projections") ReplacingEnumeratedTrace() { kind = GRLCU_TRACE_REPLACING; } };
struct AccumulatingEnumeratedTrace : Trace { TYPEINFO("trace/enumerated/accumulating", "Accumulating eligibility trace using a queue of projections") AccumulatingEnumeratedTrace() { kind = GRLCU_TRACE_ACCUMULATING; } };

class CudaTileCodingProjector;
class CudaLinearRepresentation;
class CudaSARSAPredictor;

struct Policy : Configurable { Projector *projector = nullptr; Representation *representation = nullptr; };
struct QPolicy : Policy
{
  TYPEINFO("mapping/policy/discrete/value/q", "Q-value based policy")
  UniformDiscretizer *discretizer = nullptr; Sampler *sampler = nullptr;
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};
/// ActionPolicy (policies/action.cpp:39-97): a = clamp(pi(s) + N(0, sigma)); only theta = 1 (pure Gaussian
/// noise), no decay and no renormalisation run on the device.
struct ActionPolicy : Policy
{
  TYPEINFO("mapping/policy/action", "Policy based on a direct action representation")
  double sigma = 0; Vector min, max;
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};

struct Agent : Configurable { Policy *policy = nullptr; };
struct TDAgent : Agent
{
  TYPEINFO("agent/td", "Agent that learns from observed state transitions")
  Predictor *predictor = nullptr;
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};
struct FixedAgent : Agent
{
  TYPEINFO("agent/fixed", "Fixed-policy agent")
  void request(ConfigurationRequest *config) override;
  void configure(Configuration &config) override;
};

// ---- the four CUDA types ----------------------------------------------------------------------------------
/// Shared device context of one experiment: the four objects contribute their parameters, the first
/// call that needs the device creates it (hard part 8).
struct CudaPath
{
  grlcu_config cfg;
  grlcu_ctx *ctx = nullptr;              ///< the first context (global agent ids start here): the one-agent role interfaces use it
  std::vector<grlcu_ctx *> ctxs;         ///< one context per device, contiguous ranges of global agent ids
  grlcu_group *group = nullptr;          ///< drives the contexts together; the one collective (NCCL all-gather of results)
  std::vector<int> devices;              ///< CUDA device ordinals; empty: {device}
  int agents = 1, device = 0;
  long long first_agent_id = 0;
  bool have_env = false, have_projector = false, have_representation = false, have_predictor = false;
  CudaPath();
  ~CudaPath();
  grlcu_ctx *context();                  ///< creates the contexts on first use; throws grl::Exception
  grlcu_ctx *contextOf(int agent, int *local);   ///< context and index within it of agent `agent` of this run
  int perContext() { context(); return agents / (int)ctxs.size(); }
  void release();
  static void check(int rc);             ///< non-zero C-ABI return -> grl::Exception(grlcu_last_error())
};

class CudaBatchedEnvironment : public Environment
{
  public:
    TYPEINFO("environment/model/cuda_batched", "Modeled environment stepped for many agents at once on a B200")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    void reconfigure(const Configuration &config) override;
    void start(int test, Observation *obs) override;
    double step(const Action &action, Observation *obs, double *reward, int *terminal) override;
    void contribute(CudaPath *path) const;
    void attach(CudaPath *path) { path_ = path; }
  private:
    DynamicalModel *model_ = nullptr; Task *task_ = nullptr; int discrete_time_ = 1;
    CudaPath *path_ = nullptr; Vector state_;
};

class CudaTileCodingProjector : public Projector
{
  public:
    TYPEINFO("projector/tile_coding/cuda", "Hashed tile coding projector evaluated on a B200")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    ProjectionPtr project(const Vector &in) const override;
    void contribute(CudaPath *path) const;
    void attach(CudaPath *path) { path_ = path; }
    int tilings() const { return tilings_; }
    int memory() const { return memory_; }
    const Vector &resolution() const { return resolution_; }
    const Vector &wrapping() const { return wrapping_; }
  private:
    int tilings_ = 16, memory_ = 8 * 1024 * 1024, safe_ = 0; Vector resolution_, wrapping_;
    CudaPath *path_ = nullptr;
};

class CudaLinearRepresentation : public Representation
{
  public:
    TYPEINFO("representation/linear/cuda", "Per-agent hashed linear-in-parameters representation resident in HBM")
    void request(const std::string &role, ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    void reconfigure(const Configuration &config) override;       ///< action = reset | save | load
    double read(const ProjectionPtr &projection, Vector *result) const override;
    void write(const ProjectionPtr projection, const Vector &target, const Vector &alpha) override;
    void update(const ProjectionPtr projection, const Vector &delta) override;
    void contribute(CudaPath *path) const;
    /// actor: this object is the actor-critic policy's action representation (second table of the context)
    void attach(CudaPath *path, bool actor = false) { path_ = path; actor_ = actor; }
    int memory() const { return memory_; }
    const Vector &outputMin() const { return output_min_; }
    const Vector &outputMax() const { return output_max_; }
    double initMin() const { return init_min_[0]; }
    double initMax() const { return init_max_[0]; }
  private:
    Vector init_min_, init_max_, output_min_, output_max_; int memory_ = 8 * 1024 * 1024, outputs_ = 1, limit_ = 1;
    CudaPath *path_ = nullptr; bool actor_ = false;
};

class CudaSARSAPredictor : public Predictor
{
  public:
    TYPEINFO("predictor/sarsa/cuda", "SARSA(lambda) value function predictor fused with the environment step on a B200")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    void reconfigure(const Configuration &config) override;
    virtual void contribute(CudaPath *path) const;
    Projector *projector() const { return projector_; }
    Representation *representation() const { return representation_; }
  protected:
    double alpha_ = 0.2, gamma_ = 0.97, lambda_ = 0.65;
    Projector *projector_ = nullptr; Representation *representation_ = nullptr; Trace *trace_ = nullptr;
};

/// predictor/q/cuda (stands in for predictor/critic/q, predictors/advantage.cpp:35-110): Q-learning; the SARSA
/// kernels with the target bootstrapping from max_a Q(s', a).
class CudaQPredictor : public CudaSARSAPredictor
{
  public:
    TYPEINFO("predictor/q/cuda", "Q-learning off-policy value function predictor fused with the environment step on a B200")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    void contribute(CudaPath *path) const override;
    UniformDiscretizer *discretizer() const { return discretizer_; }
  private:
    UniformDiscretizer *discretizer_ = nullptr;
};

/// predictor/expected_sarsa/cuda (stands in for predictor/critic/expected_sarsa, predictors/sarsa.cpp:134-194):
/// the target bootstraps from the value of s' under the target policy (greedy or epsilon-greedy sampler).
class CudaExpectedSARSAPredictor : public CudaSARSAPredictor
{
  public:
    TYPEINFO("predictor/expected_sarsa/cuda", "Expected SARSA value function predictor fused with the environment step on a B200")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    void contribute(CudaPath *path) const override;
    QPolicy *targetPolicy() const { return policy_; }
  private:
    QPolicy *policy_ = nullptr;
};

/// predictor/critic/td/cuda (stands in for predictor/critic/td, predictors/td.cpp:35-66): TD(lambda) on V(s);
/// its type starts with `predictor/critic`, which is what predictor/ac/action asks for (ac.cpp:47).
class CudaTDPredictor : public Predictor
{
  public:
    TYPEINFO("predictor/critic/td/cuda", "TD(lambda) state-value critic fused with the environment step on a B200")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    void contribute(CudaPath *path) const;
    Projector *projector() const { return projector_; }
    Representation *representation() const { return representation_; }
  private:
    double alpha_ = 0.2, gamma_ = 0.97, lambda_ = 0.65;
    Projector *projector_ = nullptr; Representation *representation_ = nullptr; Trace *trace_ = nullptr;
};

/// predictor/ac/action/cuda (stands in for predictor/ac/action, predictors/ac.cpp:35-70): proportional actor
/// update driven by the critic's TD error.
class CudaACPredictor : public Predictor
{
  public:
    TYPEINFO("predictor/ac/action/cuda", "Actor-critic predictor for direct action policies on a B200")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    void contribute(CudaPath *path) const;
    Projector *projector() const { return projector_; }
    Representation *representation() const { return representation_; }
    CudaTDPredictor *critic() const { return critic_; }
  private:
    double alpha_ = 0.01;
    Projector *projector_ = nullptr; Representation *representation_ = nullptr; CudaTDPredictor *critic_ = nullptr;
};

class OnlineLearningExperiment : public Experiment
{
  public:
    TYPEINFO("experiment/online_learning", "Interactive learning experiment, batched over `agents` independent agents")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    void reconfigure(const Configuration &config) override;
    std::vector<double> run() override;
    /// The C-ABI configuration this experiment resolves to (no device needed): used by --check and tests.
    const grlcu_config &resolved() { assemble(); return path_.cfg; }
    int agents() const { return agents_; }
    void setAgents(int n) { agents_ = n; assembled_ = false; }   ///< experiment/multi: instances on the agents axis
    Agent *agent() const { return agent_; }
    double collectiveMs() const { return collective_ms_total_; }   ///< device time spent in the NCCL all-gathers of run()
    CudaPath &cudaPath() { return path_; }
    unsigned long long seed = 1;
    std::string identity;
  private:
    void assemble();
    int runs_ = 1, run_offset_ = 0, trials_ = 0, steps_ = 0, rate_ = 0, test_interval_ = -1, test_trials_ = 1;
    int agents_ = 1, steps_per_launch_ = 100, device_ = -1, checkpoint_agents_ = 0;
    Vector devices_;                                       // CUDA devices the agents are sharded over (empty: see configure)
    std::string output_, load_file_, save_every_ = "never";
    Environment *environment_ = nullptr; Agent *agent_ = nullptr, *test_agent_ = nullptr;
    CudaPath path_;
    Vector sweep_alpha_, sweep_lambda_, sweep_epsilon_;   // hyper-parameter grid over the agents axis (MultiExperiment / grlo)
    std::vector<Representation *> checkpoint_reps_;     // tables written by save_every / read by load_file
    bool assembled_ = false;
    double collective_ms_total_ = 0;
};

/// experiment/multi (experiments/multi.cpp:38-75): the reference clones the prototype experiment `instances` times and
/// runs the clones on threads.  On the batched path the instances become agents: the prototype runs once with
/// instances x agents independent agents (instance i owns the agent ids [i*agents, (i+1)*agents) of each run).  That is
/// the same computation only when every clone would own its policy, i.e. when the policy's tables are configured
/// INSIDE the prototype; the reference's own example (cfg/pendulum/multi_sarsa_tc.yaml) points every clone at one
/// top-level policy, so that its threads update one shared table without synchronisation -- a different algorithm,
/// which is rejected here with an explanation instead of being silently replaced.
class MultiExperiment : public Experiment
{
  public:
    TYPEINFO("experiment/multi", "Run multiple experiments in parallel (as independent agents on the GPU)")
    void request(ConfigurationRequest *config) override;
    void configure(Configuration &config) override;
    std::vector<double> run() override;
    OnlineLearningExperiment *prototype() const { return prototype_; }
    int instances() const { return instances_; }
  private:
    int instances_ = 4;
    OnlineLearningExperiment *prototype_ = nullptr;
};

}  // namespace grl
