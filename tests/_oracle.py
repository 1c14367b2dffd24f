"""ctypes binding of the CPU parity oracle (oracle/liborc.so).

Test infrastructure only: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product never imports it.
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_DIR = os.path.join(os.path.dirname(HERE), "oracle")

MAX_DIMS, MAX_ACTIONS, MAX_TILINGS, MAX_STATE, MAX_TRACE = 8, 8, 32, 7, 64
INVALID = 0xFFFFFFFF

DYN_PENDULUM, DYN_CART_POLE, DYN_CART_DOUBLE_POLE = 0, 1, 2
TRACE_NONE, TRACE_REPLACING, TRACE_ACCUMULATING = 0, 1, 2
ALGO_SARSA, ALGO_AC, ALGO_Q, ALGO_EXPECTED_SARSA = 0, 1, 2, 3
SAMPLER_EPSILON_GREEDY, SAMPLER_SOFTMAX = 0, 1
RNG_PHILOX, RNG_GLIBC = 0, 1


class Config(C.Structure):
    """Mirror of orc_config / grlcu_config (same field order in both headers)."""
    _fields_ = [
        ("dynamics", C.c_int32), ("integration_steps", C.c_int32),
        ("control_step", C.c_double), ("timeout", C.c_double), ("randomization", C.c_double),
        ("tilings", C.c_int32), ("memory", C.c_int32), ("obs_dims", C.c_int32), ("n_actions", C.c_int32),
        ("resolution", C.c_double * MAX_DIMS), ("wrapping", C.c_double * MAX_DIMS),
        ("actions", C.c_double * MAX_ACTIONS),
        ("alpha", C.c_double), ("gamma", C.c_double), ("lambda_", C.c_double), ("epsilon", C.c_double),
        ("trace", C.c_int32), ("test_interval", C.c_int32),
        ("init_min", C.c_double), ("init_max", C.c_double),
        ("seed", C.c_uint64), ("rng", C.c_int32), ("algo", C.c_int32),
        ("sigma", C.c_double), ("actor_alpha", C.c_double),
        ("action_min", C.c_double), ("action_max", C.c_double),
        ("actor_resolution", C.c_double * MAX_DIMS), ("actor_wrapping", C.c_double * MAX_DIMS),
        ("target_epsilon", C.c_double),
        ("end_stop_penalty", C.c_int32), ("sampler", C.c_int32), ("softmax_tau", C.c_double),
        ("safe", C.c_int32),
    ]

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        if "end_stop_penalty" not in kw:
            self.end_stop_penalty = 1     # the reference's default (cart_pole.h:75)


class Record(C.Structure):
    _fields_ = [
        ("started", C.c_int32), ("start_draw", C.c_double), ("start_state", C.c_double * MAX_STATE),
        ("start_u", C.c_double), ("start_r_explore", C.c_uint32), ("start_r_tie", C.c_uint32),
        ("start_action_index", C.c_int32), ("start_q", C.c_double * MAX_ACTIONS),
        ("noise", C.c_double), ("start_noise", C.c_double),
        ("state", C.c_double * MAX_STATE), ("action", C.c_double), ("next", C.c_double * MAX_STATE),
        ("obs", C.c_double * MAX_DIMS), ("reward", C.c_double), ("terminal", C.c_int32), ("test", C.c_int32),
        ("u", C.c_double), ("r_explore", C.c_uint32), ("r_tie", C.c_uint32), ("explored", C.c_int32),
        ("hash", C.c_uint32 * (MAX_ACTIONS * MAX_TILINGS)), ("index", C.c_uint32 * (MAX_ACTIONS * MAX_TILINGS)),
        ("q", C.c_double * MAX_ACTIONS), ("action_index", C.c_int32),
        ("q_prev", C.c_double), ("target", C.c_double), ("delta", C.c_double),
        ("swept", C.c_int32), ("swept_entries", C.c_int32),
        ("trace_len", C.c_int32), ("trace_decay", C.c_double * MAX_TRACE),
        ("trace_index", C.c_uint32 * (MAX_TRACE * MAX_TILINGS)),
        ("episode_done", C.c_int32), ("episode_return", C.c_double),
    ]


def pendulum_config(memory=65536, **kw):
    """cfg/pendulum/sarsa_tc.yaml (reference) with a per-agent table size."""
    c = Config()
    c.dynamics, c.integration_steps, c.control_step = DYN_PENDULUM, 5, 0.03
    c.timeout, c.randomization = 2.99, 0.0
    c.tilings, c.memory, c.obs_dims, c.n_actions = 16, memory, 2, 3
    for i, v in enumerate([0.31415, 3.1415, 3]):
        c.resolution[i] = v
    for i, v in enumerate([6.283, 0, 0]):
        c.wrapping[i] = v
    for i, v in enumerate([-3.0, 0.0, 3.0]):
        c.actions[i] = v
    c.alpha, c.gamma, c.lambda_, c.epsilon = 0.2, 0.97, 0.65, 0.05
    c.trace, c.test_interval = TRACE_REPLACING, 10
    c.init_min, c.init_max, c.seed = 0.0, 1.0, 1
    c.rng, c.algo = RNG_PHILOX, ALGO_SARSA
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def cart_pole_config(memory=262144, **kw):
    """cfg/cart_pole/sarsa.yaml (reference) with a per-agent table size."""
    c = Config()
    c.dynamics, c.integration_steps, c.control_step = DYN_CART_POLE, 5, 0.05
    c.timeout, c.randomization = 9.99, 0.0
    c.tilings, c.memory, c.obs_dims, c.n_actions = 16, memory, 4, 3
    for i, v in enumerate([2.5, 0.157075, 2.5, 1.57075, 15]):
        c.resolution[i] = v
    for i, v in enumerate([0, 6.283, 0, 0, 0]):
        c.wrapping[i] = v
    for i, v in enumerate([-15.0, 0.0, 15.0]):
        c.actions[i] = v
    c.alpha, c.gamma, c.lambda_, c.epsilon = 0.2, 0.97, 0.65, 0.05
    c.trace, c.test_interval = TRACE_REPLACING, 10
    c.init_min, c.init_max, c.seed = 0.0, 1.0, 1
    c.rng, c.algo = RNG_PHILOX, ALGO_SARSA
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def cart_double_pole_config(memory=131072, **kw):
    """cfg/cart_double_pole/sarsa.yaml (reference) with a per-agent table size."""
    c = Config()
    c.dynamics, c.integration_steps, c.control_step = DYN_CART_DOUBLE_POLE, 5, 0.05
    c.timeout, c.randomization = 4.99, 0.0
    c.tilings, c.memory, c.obs_dims, c.n_actions = 16, memory, 6, 3
    for i, v in enumerate([2.5, 0.0785375, 0.0785375, 2.5, 0.785375, 0.785375, 20]):
        c.resolution[i] = v
    for i, v in enumerate([-20.0, 0.0, 20.0]):
        c.actions[i] = v
    c.alpha, c.gamma, c.lambda_, c.epsilon = 0.2, 0.99, 0.92, 0.02
    c.trace, c.test_interval = TRACE_REPLACING, 10
    c.init_min, c.init_max, c.seed = 20.0, 21.0, 1
    c.rng, c.algo = RNG_PHILOX, ALGO_SARSA
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def cart_double_pole_ac_config(memory=131072, **kw):
    """cfg/cart_double_pole/ac.yaml (reference): actor-critic with two tile-coded tables per agent."""
    c = Config()
    c.dynamics, c.integration_steps, c.control_step = DYN_CART_DOUBLE_POLE, 5, 0.05
    c.timeout, c.randomization = 4.99, 0.0
    c.tilings, c.memory, c.obs_dims, c.n_actions = 16, memory, 6, 0
    for i, v in enumerate([2.5, 0.0785375, 0.0785375, 2.5, 0.785375, 0.785375]):
        c.resolution[i] = v
        c.actor_resolution[i] = v
    c.alpha, c.gamma, c.lambda_, c.epsilon = 0.2, 0.97, 0.65, 0.0
    c.trace, c.test_interval = TRACE_REPLACING, 10
    c.init_min, c.init_max, c.seed = 0.0, 1.0, 1
    c.rng, c.algo = RNG_PHILOX, ALGO_AC
    c.sigma, c.actor_alpha, c.action_min, c.action_max = 5.0, 0.01, -20.0, 20.0
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def _ac_config(dynamics, control_step, timeout_, memory, res, wrap, sigma_, amax, **kw):
    c = Config()
    c.dynamics, c.integration_steps, c.control_step = dynamics, 5, control_step
    c.timeout, c.randomization = timeout_, 0.0
    c.tilings, c.memory, c.obs_dims, c.n_actions = 16, memory, len(res), 0
    for i, (r, w) in enumerate(zip(res, wrap)):
        c.resolution[i] = c.actor_resolution[i] = r
        c.wrapping[i] = c.actor_wrapping[i] = w
    c.alpha, c.gamma, c.lambda_, c.epsilon = 0.2, 0.97, 0.65, 0.0
    c.trace, c.test_interval = TRACE_REPLACING, 10
    c.init_min, c.init_max, c.seed = 0.0, 1.0, 1
    c.rng, c.algo = RNG_PHILOX, ALGO_AC
    c.sigma, c.actor_alpha, c.action_min, c.action_max = sigma_, 0.01, -amax, amax
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def pendulum_ac_config(memory=65536, **kw):
    """cfg/pendulum/ac_tc.yaml (reference)."""
    return _ac_config(DYN_PENDULUM, 0.03, 2.99, memory, [0.31415, 3.1415], [6.283, 0], 1.0, 3.0, **kw)


def cart_pole_ac_config(memory=262144, **kw):
    """cfg/cart_pole/ac_tc.yaml (reference): end_stop_penalty 0."""
    kw.setdefault("end_stop_penalty", 0)
    return _ac_config(DYN_CART_POLE, 0.05, 9.99, memory, [2.5, 0.157075, 2.5, 1.57075], [0, 6.283, 0, 0], 5.0, 15.0, **kw)


# ---- YAML for the reference's own deployer (oracle/_ref/grld_ref: tests/test_oracle_vs_grld_ref.py, bench.py --impl reference)
GRLD_TYPES = {DYN_PENDULUM: ("dynamics/pendulum", "task/pendulum/swingup"),
       DYN_CART_POLE: ("dynamics/cart_pole", "task/cart_pole/swingup"),
       DYN_CART_DOUBLE_POLE: ("dynamics/cart_double_pole", "task/cart_double_pole/balancing")}


def _vec(a, n):
    return "[" + ", ".join(repr(float(a[i])) for i in range(n)) + "]"


def _environment(c):
    dyn, task = GRLD_TYPES[c.dynamics]
    extra = ""
    if c.dynamics == DYN_CART_POLE:
        extra = ", end_stop_penalty: %d" % c.end_stop_penalty
    # (task/cart_pole/swingup declares `randomization` as an integer, task/pendulum/swingup as a real number)
    rand = "" if c.dynamics == DYN_CART_DOUBLE_POLE and c.randomization == 0 else ", randomization: %s" % (
        "%d" % int(c.randomization) if c.dynamics == DYN_CART_POLE else repr(c.randomization))
    return """  environment:
    type: environment/modeled
    model:
      type: model/dynamical
      control_step: %r
      integration_steps: %d
      dynamics: {type: %s}
    task: {type: %s, timeout: %r%s%s}
""" % (c.control_step, c.integration_steps, dyn, task, c.timeout, rand, extra)


def _trace(c):
    return {TRACE_REPLACING: "{type: trace/enumerated/replacing}", TRACE_ACCUMULATING: "{type: trace/enumerated/accumulating}",
            TRACE_NONE: "0"}[c.trace]


def grld_yaml(c, trials, output):
    """A configuration file of the shape of the reference's cfg/pendulum/sarsa_tc.yaml / cfg/cart_double_pole/ac.yaml
    (visualisation left out), generated from the oracle's configuration so that both sides always run the same thing."""
    head = """experiment:
  type: experiment/online_learning
  runs: 1
  trials: %d
  steps: 0
  rate: 0
  test_interval: %d
  output: %s
%s""" % (trials, c.test_interval, output, _environment(c))
    if c.algo == ALGO_AC:
        n = c.obs_dims
        return head + """  agent:
    type: agent/td
    policy:
      type: mapping/policy/action
      sigma: [%r]
      output_min: experiment/environment/task/action_min
      output_max: experiment/environment/task/action_max
      projector:
        type: projector/tile_coding
        tilings: %d
        memory: %d
        safe: 0
        resolution: %s
        wrapping: %s
      representation:
        type: representation/parameterized/linear
        init_min: [%r]
        init_max: [%r]
        memory: experiment/agent/policy/projector/memory
        outputs: experiment/environment/task/action_dims
        output_min: experiment/environment/task/action_min
        output_max: experiment/environment/task/action_max
    predictor:
      type: predictor/ac/action
      alpha: %r
      update_method: proportional
      step_limit: []
      critic:
        type: predictor/critic/td
        alpha: %r
        gamma: %r
        lambda: %r
        projector: {type: projector/tile_coding, tilings: %d, memory: %d, safe: 0, resolution: %s, wrapping: %s}
        representation:
          type: representation/parameterized/linear
          init_min: [%r]
          init_max: [%r]
          memory: experiment/agent/predictor/critic/projector/memory
          outputs: 1
          output_min: []
          output_max: []
        trace: %s
      projector: experiment/agent/policy/projector
      representation: experiment/agent/policy/representation
  test_agent:
    type: agent/fixed
    policy:
      type: mapping/policy/action
      sigma: []
      output_min: experiment/environment/task/action_min
      output_max: experiment/environment/task/action_max
      projector: experiment/agent/policy/projector
      representation: experiment/agent/policy/representation
  save_every: never
""" % (c.sigma, c.tilings, c.memory, _vec(c.actor_resolution, n), _vec(c.actor_wrapping, n), c.init_min, c.init_max,
       c.actor_alpha, c.alpha, c.gamma, c.lambda_, c.tilings, c.memory, _vec(c.resolution, n), _vec(c.wrapping, n),
       c.init_min, c.init_max, _trace(c))
    n = c.obs_dims + 1
    sampler = ("{type: sampler/softmax, tau: %r}" % c.softmax_tau if c.sampler == SAMPLER_SOFTMAX
               else "{type: sampler/epsilon_greedy, epsilon: %r}" % c.epsilon)
    predictor = {ALGO_SARSA: "predictor/critic/sarsa", ALGO_Q: "predictor/critic/q",
                 ALGO_EXPECTED_SARSA: "predictor/critic/expected_sarsa"}[c.algo]
    extra = "      discretizer: experiment/agent/policy/discretizer\n" if c.algo == ALGO_Q else ""
    if c.algo == ALGO_EXPECTED_SARSA:
        # the target policy is the agent's own (epsilon-greedy) policy object: no further sampler is constructed, so the
        # random streams stay where the oracle's glibc mode puts them
        assert c.target_epsilon == c.epsilon and c.sampler == SAMPLER_EPSILON_GREEDY
        extra = "      policy: experiment/agent/policy\n"
    return head + """  agent:
    type: agent/td
    policy:
      type: mapping/policy/discrete/value/q
      discretizer:
        type: discretizer/uniform
        min: experiment/environment/task/action_min
        max: experiment/environment/task/action_max
        steps: [%d]
      projector:
        type: projector/tile_coding
        tilings: %d
        memory: %d
        safe: %d
        resolution: %s
        wrapping: %s
      representation:
        type: representation/parameterized/linear
        init_min: [%r]
        init_max: [%r]
        memory: experiment/agent/policy/projector/memory
        outputs: 1
        output_min: []
        output_max: []
      sampler: %s
    predictor:
      type: %s
      alpha: %r
      gamma: %r
      lambda: %r
%s      projector: experiment/agent/policy/projector
      representation: experiment/agent/policy/representation
      trace: %s
  test_agent:
    type: agent/fixed
    policy:
      type: mapping/policy/discrete/value/q
      discretizer: experiment/agent/policy/discretizer
      projector: experiment/agent/policy/projector
      representation: experiment/agent/policy/representation
      sampler: {type: sampler/greedy}
  save_every: never
""" % (c.n_actions, c.tilings, c.memory, c.safe, _vec(c.resolution, n), _vec(c.wrapping, n), c.init_min, c.init_max,
       sampler, predictor, c.alpha, c.gamma, c.lambda_, extra, _trace(c))



_lib = None


def build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(ORACLE_DIR, "liborc.so")
    if not os.path.exists(path):
        build()
    L = C.CDLL(path)
    dp, u32p, i32p = C.POINTER(C.c_double), C.POINTER(C.c_uint32), C.POINTER(C.c_int32)
    L.orc_hash.restype = C.c_uint32
    L.orc_hash.argtypes = [i32p, C.c_uint32]
    L.orc_murmur2_bytes.restype = C.c_uint32
    L.orc_murmur2_bytes.argtypes = [C.c_void_p, C.c_int, C.c_uint32]
    L.orc_project.restype = None
    L.orc_project.argtypes = [dp, dp, C.c_int, C.c_int, C.c_int32, dp, u32p, u32p]
    L.orc_eom.restype = None
    L.orc_eom.argtypes = [C.c_int, dp, C.c_double, dp]
    L.orc_rk4.restype = None
    L.orc_rk4.argtypes = [C.c_int, C.c_double, C.c_int, dp, C.c_double, dp]
    L.orc_state_dims.restype = C.c_int
    L.orc_state_dims.argtypes = [C.c_int]
    L.orc_env_step.restype = None
    L.orc_env_step.argtypes = [C.POINTER(Config), dp, C.c_double, dp, dp, dp, i32p]
    L.orc_philox.restype = None
    L.orc_philox.argtypes = [C.c_uint32] * 6 + [u32p]
    L.orc_create.restype = C.c_void_p
    L.orc_create.argtypes = [C.POINTER(Config), C.c_uint64]
    L.orc_destroy.restype = None
    L.orc_destroy.argtypes = [C.c_void_p]
    L.orc_set_hyper.restype = None
    L.orc_set_hyper.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double]
    L.orc_glibc_seed.restype = None
    L.orc_glibc_seed.argtypes = [C.c_long]
    L.orc_table.restype = dp
    L.orc_table.argtypes = [C.c_void_p]
    L.orc_actor_table.restype = dp
    L.orc_actor_table.argtypes = [C.c_void_p]
    L.orc_step.restype = C.c_int
    L.orc_step.argtypes = [C.c_void_p, C.POINTER(Record)]
    L.orc_run.restype = C.c_uint64
    L.orc_run.argtypes = [C.c_void_p, C.c_uint64]
    L.orc_stats.restype = None
    L.orc_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                            C.POINTER(C.c_uint64), dp, dp]
    _lib = L
    return L


class OracleBatch:
    """N independent oracle agents (global ids first_id..), stepped in lock-step, with the records
    of a step gathered into SoA numpy arrays shaped like grlcu_replay_in/out."""

    def __init__(self, cfg, n, first_id=0):
        import numpy as np
        self.np = np
        self.L, self.cfg, self.n = lib(), cfg, n
        self.S = self.L.orc_state_dims(cfg.dynamics)
        self.agents = [self.L.orc_create(C.byref(cfg), first_id + i) for i in range(n)]
        self.recs = [Record() for _ in range(n)]

    def close(self):
        for a in self.agents:
            self.L.orc_destroy(a)
        self.agents = []

    def table(self, i):
        return self.np.ctypeslib.as_array(self.L.orc_table(self.agents[i]), shape=(self.cfg.memory,))

    def actor_table(self, i):
        return self.np.ctypeslib.as_array(self.L.orc_actor_table(self.agents[i]), shape=(self.cfg.memory,))

    def set_hyper(self, alpha, lambda_, epsilon):
        for i, a in enumerate(self.agents):
            self.L.orc_set_hyper(a, float(alpha[i]), float(lambda_[i]), float(epsilon[i]))

    def step(self):
        np, n, S, OBS, T, A = self.np, self.n, self.S, self.cfg.obs_dims, self.cfg.tilings, self.cfg.n_actions
        ac = self.cfg.algo == ALGO_AC
        if ac:
            A = 4          # AC record: q rows = pi(s'), V(s'), u, a'; hash/index rows = actor, critic projection
        for a, r in zip(self.agents, self.recs):
            self.L.orc_step(a, C.byref(r))
        R = self.recs
        f64 = lambda name: np.array([getattr(r, name) for r in R], dtype=np.float64)
        i32 = lambda name: np.array([getattr(r, name) for r in R], dtype=np.int32)
        u32 = lambda name: np.array([getattr(r, name) for r in R], dtype=np.uint32)
        vec = lambda name, k: np.array([list(getattr(r, name))[:k] for r in R], dtype=np.float64).T.copy()
        d = dict(started=i32("started"), start_state=vec("start_state", S), start_u=f64("start_u"),
                 start_r_explore=u32("start_r_explore"), start_r_tie=u32("start_r_tie"),
                 state=vec("state", S), action=f64("action"), test=i32("test"),
                 obs=vec("obs", OBS), reward=f64("reward"), terminal=i32("terminal"),
                 u=f64("u"), r_explore=u32("r_explore"), r_tie=u32("r_tie"),
                 next=vec("next", S), q=vec("q", A), action_index=i32("action_index"),
                 start_q=vec("start_q", A), start_action_index=i32("start_action_index"),
                 q_prev=f64("q_prev"), target=f64("target"), delta=f64("delta"),
                 swept=i32("swept"), trace_len=i32("trace_len"),
                 episode_done=i32("episode_done"), episode_return=f64("episode_return"))
        HA = 2 if ac else A
        hs = np.zeros((HA, T, n), np.uint32)
        ix = np.zeros((HA, T, n), np.uint32)
        for i, r in enumerate(R):
            h = np.ctypeslib.as_array(r.hash).reshape(MAX_ACTIONS, MAX_TILINGS)
            x = np.ctypeslib.as_array(r.index).reshape(MAX_ACTIONS, MAX_TILINGS)
            hs[:, :, i] = h[:HA, :T]
            ix[:, :, i] = x[:HA, :T]
        d["hash"], d["index"] = hs, ix
        if ac:
            # the draw that is fed in replay mode is the Gaussian exploration term itself
            d["noise"], d["start_noise"] = f64("noise"), f64("start_noise")
            d["u"], d["start_u"] = d["noise"], d["start_noise"]
        return d

    def trace(self, cap):
        """[cap][T][n] newest first, INVALID beyond trace_len (matches grlcu's dump)."""
        np, T = self.np, self.cfg.tilings
        out = np.full((cap, T, self.n), INVALID, np.uint32)
        for i, r in enumerate(self.recs):
            t = np.ctypeslib.as_array(r.trace_index).reshape(MAX_TRACE, MAX_TILINGS)
            k = min(r.trace_len, cap)
            out[:k, :, i] = t[:k, :T]
        return out
