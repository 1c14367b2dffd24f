"""ctypes binding of libgrlcu.so (include/grlcu.h).  No fallback: loading or creating a context
fails loudly when the CUDA library or a GPU is missing."""
import ctypes as C
import os

import numpy as np

from . import build as _build

MAX_DIMS, MAX_ACTIONS, MAX_STATE = 8, 8, 7
DYN_PENDULUM, DYN_CART_POLE, DYN_CART_DOUBLE_POLE = 0, 1, 2
TRACE_NONE, TRACE_REPLACING, TRACE_ACCUMULATING = 0, 1, 2
ALGO_SARSA, ALGO_AC, ALGO_Q, ALGO_EXPECTED_SARSA = 0, 1, 2, 3
SAMPLER_EPSILON_GREEDY, SAMPLER_SOFTMAX = 0, 1
RESET_TABLES, RESET_AGENTS, RESET_STATS = 1, 2, 4
INVALID = 0xFFFFFFFF


class GrlcuError(RuntimeError):
    pass


class Config(C.Structure):
    """grlcu_config (field names follow GRL's YAML keys; `lambda_` = lambda)."""
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


def _fill(c, **kw):
    for k, v in kw.items():
        if isinstance(v, (list, tuple)):
            arr = getattr(c, k)
            for i, x in enumerate(v):
                arr[i] = x
        else:
            setattr(c, k, v)
    return c


def pendulum_config(memory=65536, **kw):
    """cfg/pendulum/sarsa_tc.yaml of the reference, with a per-agent table size."""
    c = Config()
    _fill(c, dynamics=DYN_PENDULUM, integration_steps=5, control_step=0.03, timeout=2.99, randomization=0.0,
          tilings=16, memory=memory, obs_dims=2, n_actions=3,
          resolution=[0.31415, 3.1415, 3], wrapping=[6.283, 0, 0], actions=[-3.0, 0.0, 3.0],
          alpha=0.2, gamma=0.97, lambda_=0.65, epsilon=0.05, trace=TRACE_REPLACING, test_interval=10,
          init_min=0.0, init_max=1.0, seed=1, algo=ALGO_SARSA)
    return _fill(c, **kw)


def cart_pole_config(memory=262144, **kw):
    """cfg/cart_pole/sarsa.yaml of the reference, with a per-agent table size."""
    c = Config()
    _fill(c, dynamics=DYN_CART_POLE, integration_steps=5, control_step=0.05, timeout=9.99, randomization=0.0,
          tilings=16, memory=memory, obs_dims=4, n_actions=3,
          resolution=[2.5, 0.157075, 2.5, 1.57075, 15], wrapping=[0, 6.283, 0, 0, 0],
          actions=[-15.0, 0.0, 15.0],
          alpha=0.2, gamma=0.97, lambda_=0.65, epsilon=0.05, trace=TRACE_REPLACING, test_interval=10,
          init_min=0.0, init_max=1.0, seed=1, algo=ALGO_SARSA)
    return _fill(c, **kw)


def cart_double_pole_config(memory=131072, **kw):
    """cfg/cart_double_pole/sarsa.yaml of the reference (SARSA(lambda) tile coding on the cart-double-pole balancing task:
    6-D observations, 3 actions, gamma 0.99, lambda 0.92 => a 51-entry trace ring, optimistic init 20..21)."""
    c = Config()
    _fill(c, dynamics=DYN_CART_DOUBLE_POLE, integration_steps=5, control_step=0.05, timeout=4.99, randomization=0.0,
          tilings=16, memory=memory, obs_dims=6, n_actions=3,
          resolution=[2.5, 0.0785375, 0.0785375, 2.5, 0.785375, 0.785375, 20], wrapping=[0] * 7,
          actions=[-20.0, 0.0, 20.0],
          alpha=0.2, gamma=0.99, lambda_=0.92, epsilon=0.02, trace=TRACE_REPLACING, test_interval=10,
          init_min=20.0, init_max=21.0, seed=1, algo=ALGO_SARSA)
    return _fill(c, **kw)


def cart_double_pole_ac_config(memory=131072, **kw):
    """cfg/cart_double_pole/ac.yaml of the reference (actor-critic, two tile-coded tables per agent)."""
    c = Config()
    res = [2.5, 0.0785375, 0.0785375, 2.5, 0.785375, 0.785375]
    _fill(c, dynamics=DYN_CART_DOUBLE_POLE, integration_steps=5, control_step=0.05, timeout=4.99, randomization=0.0,
          tilings=16, memory=memory, obs_dims=6, n_actions=0, resolution=res, actor_resolution=res,
          alpha=0.2, gamma=0.97, lambda_=0.65, epsilon=0.0, trace=TRACE_REPLACING, test_interval=10,
          init_min=0.0, init_max=1.0, seed=1, algo=ALGO_AC,
          sigma=5.0, actor_alpha=0.01, action_min=-20.0, action_max=20.0)
    return _fill(c, **kw)


def pendulum_ac_config(memory=65536, **kw):
    """cfg/pendulum/ac_tc.yaml of the reference (actor-critic on the pendulum swing-up, sigma = 1, action +-3)."""
    c = Config()
    res, wrap = [0.31415, 3.1415], [6.283, 0]
    _fill(c, dynamics=DYN_PENDULUM, integration_steps=5, control_step=0.03, timeout=2.99, randomization=0.0,
          tilings=16, memory=memory, obs_dims=2, n_actions=0, resolution=res, wrapping=wrap,
          actor_resolution=res, actor_wrapping=wrap,
          alpha=0.2, gamma=0.97, lambda_=0.65, epsilon=0.0, trace=TRACE_REPLACING, test_interval=10,
          init_min=0.0, init_max=1.0, seed=1, algo=ALGO_AC,
          sigma=1.0, actor_alpha=0.01, action_min=-3.0, action_max=3.0)
    return _fill(c, **kw)


def cart_pole_ac_config(memory=262144, **kw):
    """cfg/cart_pole/ac_tc.yaml of the reference (actor-critic on the cart-pole swing-up: sigma = 5, action +-15,
    end_stop_penalty 0: the cart rides the end stop instead of ending the episode)."""
    c = Config()
    res, wrap = [2.5, 0.157075, 2.5, 1.57075], [0, 6.283, 0, 0]
    _fill(c, dynamics=DYN_CART_POLE, integration_steps=5, control_step=0.05, timeout=9.99, randomization=0.0,
          tilings=16, memory=memory, obs_dims=4, n_actions=0, resolution=res, wrapping=wrap,
          actor_resolution=res, actor_wrapping=wrap,
          alpha=0.2, gamma=0.97, lambda_=0.65, epsilon=0.0, trace=TRACE_REPLACING, test_interval=10,
          init_min=0.0, init_max=1.0, seed=1, algo=ALGO_AC,
          sigma=5.0, actor_alpha=0.01, action_min=-15.0, action_max=15.0, end_stop_penalty=0)
    return _fill(c, **kw)


_P = C.POINTER
_dp, _u32p, _i32p, _u64p = _P(C.c_double), _P(C.c_uint32), _P(C.c_int32), _P(C.c_uint64)


class ReplayIn(C.Structure):
    _fields_ = [("started", _i32p), ("start_state", _dp), ("start_u", _dp),
                ("start_r_explore", _u32p), ("start_r_tie", _u32p),
                ("state", _dp), ("action", _dp), ("test", _i32p),
                ("obs", _dp), ("reward", _dp), ("terminal", _i32p),
                ("u", _dp), ("r_explore", _u32p), ("r_tie", _u32p), ("force_action", _i32p)]


class ReplayOut(C.Structure):
    _fields_ = [("next", _dp), ("obs", _dp), ("reward", _dp), ("terminal", _i32p),
                ("hash", _u32p), ("index", _u32p), ("q", _dp), ("action_index", _i32p),
                ("start_q", _dp), ("start_action_index", _i32p),
                ("q_prev", _dp), ("target", _dp), ("delta", _dp),
                ("swept", _i32p), ("trace_len", _i32p), ("trace_index", _u32p)]


class Results(C.Structure):
    _fields_ = [("last_return", _dp), ("last_test_return", _dp), ("return_sum", _dp),
                ("steps", _u64p), ("episodes", _u32p), ("swept", _u64p), ("table_writes", _u64p)]


EXPORTS = ["grlcu_last_error", "grlcu_version", "grlcu_create", "grlcu_destroy", "grlcu_reset",
           "grlcu_set_hyper", "grlcu_project", "grlcu_model_step", "grlcu_run", "grlcu_sync",
           "grlcu_run_host", "grlcu_replay_step", "grlcu_get_results", "grlcu_get_table",
           "grlcu_set_table", "grlcu_get_state", "grlcu_trace_capacity", "grlcu_launch_count",
           "grlcu_run_kernel_time", "grlcu_task_start", "grlcu_read", "grlcu_update",
           "grlcu_get_actor_table", "grlcu_set_actor_table", "grlcu_read_actor", "grlcu_update_actor", "grlcu_probe_random_access",
           "grlcu_layout_info", "grlcu_get_pressure", "grlcu_project_noalloc", "grlcu_device_count", "grlcu_group_create", "grlcu_group_destroy", "grlcu_group_run", "grlcu_group_results"]

_lib = None


def load(build=True):
    """dlopen grl_b200/csrc/libgrlcu.so (building it in-tree first if it is stale)."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.build() if build else _build.LIB
    path = os.environ.get("GRLCU_LIB", path)      # A/B of two builds of this same library (tools/gpu_*.sh)
    if not os.path.exists(path):
        raise GrlcuError("libgrlcu.so is missing (%s): run grl_b200/build.py; there is no CPU fallback" % path)
    L = C.CDLL(path)
    vp = C.c_void_p
    L.grlcu_last_error.restype = C.c_char_p
    L.grlcu_create.argtypes = [_P(Config), C.c_int32, C.c_int64, C.c_int32, vp, _P(vp)]
    L.grlcu_destroy.argtypes = [vp]
    L.grlcu_reset.argtypes = [vp, C.c_int32]
    L.grlcu_set_hyper.argtypes = [vp, _dp, _dp, _dp]
    L.grlcu_project.argtypes = [vp, C.c_int64, _dp, _u32p, _u32p]
    L.grlcu_get_pressure.argtypes = [vp, _u32p]
    L.grlcu_project_noalloc.argtypes = [vp, C.c_int64, _dp, _u32p, _u32p]
    L.grlcu_model_step.argtypes = [vp, C.c_int64, _dp, _dp, _dp, _dp, _dp, _i32p]
    L.grlcu_run.argtypes = [vp, C.c_int64, C.c_int32]
    L.grlcu_sync.argtypes = [vp]
    L.grlcu_run_host.argtypes = [vp, _dp, C.c_int64, C.c_int32, _P(Results)]
    L.grlcu_replay_step.argtypes = [vp, _P(ReplayIn), _P(ReplayOut)]
    L.grlcu_get_results.argtypes = [vp, _P(Results)]
    L.grlcu_get_table.argtypes = [vp, C.c_int32, _dp]
    L.grlcu_set_table.argtypes = [vp, C.c_int32, _dp]
    L.grlcu_get_actor_table.argtypes = [vp, C.c_int32, _dp]
    L.grlcu_set_actor_table.argtypes = [vp, C.c_int32, _dp]
    L.grlcu_get_state.argtypes = [vp, C.c_int32, _dp, _dp, _i32p]
    L.grlcu_trace_capacity.argtypes = [vp]
    L.grlcu_launch_count.argtypes = [vp]
    L.grlcu_launch_count.restype = C.c_int64
    L.grlcu_run_kernel_time.argtypes = [vp, _dp, _P(C.c_int64)]
    L.grlcu_task_start.argtypes = [vp, C.c_int64, C.c_int32, _dp, _dp, _dp, _i32p]
    L.grlcu_read.argtypes = [vp, C.c_int32, C.c_int32, _u64p, _dp]
    L.grlcu_update.argtypes = [vp, C.c_int32, C.c_int32, _u64p, C.c_double]
    L.grlcu_probe_random_access.argtypes = [C.c_int32, C.c_uint64, _dp, _dp, _dp]
    L.grlcu_read_actor.argtypes = [vp, C.c_int32, C.c_int32, _u64p, _dp]
    L.grlcu_update_actor.argtypes = [vp, C.c_int32, C.c_int32, _u64p, C.c_double]
    L.grlcu_layout_info.argtypes = [vp, _i32p, _u32p, _u32p, _u32p]
    L.grlcu_group_create.argtypes = [_P(vp), C.c_int32, _P(vp)]
    L.grlcu_group_destroy.argtypes = [vp]
    L.grlcu_group_run.argtypes = [vp, C.c_int64]
    L.grlcu_group_results.argtypes = [vp, _P(Results), _dp]
    _lib = L
    return L


def _ptr(a, ty):
    return None if a is None else a.ctypes.data_as(ty)


def probe_random_access(device=0, footprint_bytes=1 << 33):
    """(reads/s, writes/s, rmws/s) of independent random 8-byte accesses on this device (grlcu_probe_random_access)."""
    L = load()
    r, w, m = C.c_double(), C.c_double(), C.c_double()
    rc = L.grlcu_probe_random_access(device, footprint_bytes, C.byref(r), C.byref(w), C.byref(m))
    if rc != 0:
        raise GrlcuError("grlcu error %d: %s" % (rc, L.grlcu_last_error().decode()))
    return r.value, w.value, m.value


class Batch:
    """A batch of independent agents on one GPU (one grlcu context)."""

    def __init__(self, cfg, n_agents, first_agent_id=0, device=0, stream=None):
        self.lib = load()
        self.cfg, self.n = cfg, int(n_agents)
        self.S = {0: 3, 1: 5, 2: 7}[cfg.dynamics]
        self.h = C.c_void_p()
        self._chk(self.lib.grlcu_create(C.byref(cfg), self.n, first_agent_id, device,
                                        C.c_void_p(stream or 0), C.byref(self.h)))
        self.trace_cap = self.lib.grlcu_trace_capacity(self.h)

    def _chk(self, rc):
        if rc != 0:
            raise GrlcuError("grlcu error %d: %s" % (rc, self.lib.grlcu_last_error().decode()))

    def close(self):
        if self.h:
            self.lib.grlcu_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self, what=RESET_TABLES | RESET_AGENTS | RESET_STATS):
        self._chk(self.lib.grlcu_reset(self.h, what))

    def set_hyper(self, alpha=None, lambda_=None, epsilon=None):
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float64) for a in (alpha, lambda_, epsilon)]
        self._chk(self.lib.grlcu_set_hyper(self.h, *[_ptr(a, _dp) for a in arrs]))

    def project(self, inputs):
        x = np.ascontiguousarray(inputs, dtype=np.float64)
        n, T = x.shape[0], self.cfg.tilings
        h, i = np.empty((n, T), np.uint32), np.empty((n, T), np.uint32)
        self._chk(self.lib.grlcu_project(self.h, n, _ptr(x, _dp), _ptr(h, _u32p), _ptr(i, _u32p)))
        return h, i

    def project_noalloc(self, inputs):
        """grlcu_project_noalloc: the same projection staged in the library's static device scratch (no allocation)."""
        x = np.ascontiguousarray(inputs, dtype=np.float64)
        n, T = x.shape[0], self.cfg.tilings
        h, i = np.empty((n, T), np.uint32), np.empty((n, T), np.uint32)
        self._chk(self.lib.grlcu_project_noalloc(self.h, n, _ptr(x, _dp), _ptr(h, _u32p), _ptr(i, _u32p)))
        return h, i

    def model_step(self, state, action):
        s = np.ascontiguousarray(state, dtype=np.float64)
        a = np.ascontiguousarray(action, dtype=np.float64)
        n = s.shape[0]
        nxt, obs = np.empty((n, self.S)), np.empty((n, self.cfg.obs_dims))
        rew, term = np.empty(n), np.empty(n, np.int32)
        self._chk(self.lib.grlcu_model_step(self.h, n, _ptr(s, _dp), _ptr(a, _dp), _ptr(nxt, _dp), _ptr(obs, _dp),
                                            _ptr(rew, _dp), _ptr(term, _i32p)))
        return nxt, obs, rew, term

    def task_start(self, u, test=0):
        """Task::start + observe for len(u) episodes: (state [n][S], obs [n][obs_dims], terminal [n])."""
        u = np.ascontiguousarray(u, dtype=np.float64)
        n = u.shape[0]
        st, obs, term = np.empty((n, self.S)), np.empty((n, self.cfg.obs_dims)), np.empty(n, np.int32)
        self._chk(self.lib.grlcu_task_start(self.h, n, int(test), _ptr(u, _dp), _ptr(st, _dp), _ptr(obs, _dp),
                                            _ptr(term, _i32p)))
        return st, obs, term

    def read(self, agent, indices, actor=False):
        """LinearRepresentation::read of one projection on one agent's table (actor=True: the AC actor table)."""
        idx = np.ascontiguousarray(indices, dtype=np.uint64)
        v = C.c_double()
        f = self.lib.grlcu_read_actor if actor else self.lib.grlcu_read
        self._chk(f(self.h, int(agent), idx.shape[0], _ptr(idx, _u64p), C.byref(v)))
        return v.value

    def update(self, agent, indices, delta, actor=False):
        """LinearRepresentation::update(projection, delta) on one agent's table."""
        idx = np.ascontiguousarray(indices, dtype=np.uint64)
        f = self.lib.grlcu_update_actor if actor else self.lib.grlcu_update
        self._chk(f(self.h, int(agent), idx.shape[0], _ptr(idx, _u64p), C.c_double(delta)))

    def run(self, n_steps, cohort=0):
        self._chk(self.lib.grlcu_run(self.h, n_steps, cohort))

    def sync(self):
        self._chk(self.lib.grlcu_sync(self.h))

    def _results_struct(self, bufs):
        r = Results()
        r.last_return, r.last_test_return = _ptr(bufs["last_return"], _dp), _ptr(bufs["last_test_return"], _dp)
        r.return_sum, r.steps = _ptr(bufs["return_sum"], _dp), _ptr(bufs["steps"], _u64p)
        r.episodes, r.swept = _ptr(bufs["episodes"], _u32p), _ptr(bufs["swept"], _u64p)
        r.table_writes = _ptr(bufs["table_writes"], _u64p)
        return r

    def result_buffers(self):
        n = self.n
        return dict(last_return=np.empty(n), last_test_return=np.empty(n), return_sum=np.empty(n),
                    steps=np.empty(n, np.uint64), episodes=np.empty(n, np.uint32), swept=np.empty(n, np.uint64),
                    table_writes=np.empty(n, np.uint64))

    def results(self, bufs=None):
        bufs = bufs or self.result_buffers()
        r = self._results_struct(bufs)
        self._chk(self.lib.grlcu_get_results(self.h, C.byref(r)))
        return bufs

    def run_host(self, start_states, n_steps, cohort=0, bufs=None):
        """End-to-end call with host buffers: H2D start states, n_steps per agent, D2H results."""
        bufs = bufs or self.result_buffers()
        r = self._results_struct(bufs)
        s = None if start_states is None else np.ascontiguousarray(start_states, dtype=np.float64)
        self._chk(self.lib.grlcu_run_host(self.h, _ptr(s, _dp), n_steps, cohort, C.byref(r)))
        return bufs

    def get_table(self, agent):
        out = np.empty(self.cfg.memory)
        self._chk(self.lib.grlcu_get_table(self.h, agent, _ptr(out, _dp)))
        return out

    def set_table(self, agent, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        assert v.size == self.cfg.memory
        self._chk(self.lib.grlcu_set_table(self.h, agent, _ptr(v, _dp)))

    def get_actor_table(self, agent):
        out = np.empty(self.cfg.memory)
        self._chk(self.lib.grlcu_get_actor_table(self.h, agent, _ptr(out, _dp)))
        return out

    def get_state(self, agent):
        st, act, ep = np.empty(self.S), C.c_double(), C.c_int32()
        self._chk(self.lib.grlcu_get_state(self.h, agent, _ptr(st, _dp), C.byref(act), C.byref(ep)))
        return st, act.value, ep.value

    def layout_info(self):
        """dict(tile_major, table_stride, inbox_words, shared_words): how the context stores its tables."""
        tm, ts, iw, sw = C.c_int32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._chk(self.lib.grlcu_layout_info(self.h, C.byref(tm), C.byref(ts), C.byref(iw), C.byref(sw)))
        kind = {0: "raw hash order", 1: "tile-major (slots from tile coordinates)", 2: "visit-ordered (pilot run, perm lookup)"}[tm.value]
        return dict(tile_major=bool(tm.value), kind=kind, table_stride=ts.value, inbox_words=iw.value, shared_words=sw.value)

    def pressure(self):
        """safe > 0: per-agent maximum probe length in the claim table (TileCodingProjector::pressure_)."""
        out = np.empty(self.n, np.uint32)
        self._chk(self.lib.grlcu_get_pressure(self.h, _ptr(out, _u32p)))
        return out

    def launch_count(self):
        return self.lib.grlcu_launch_count(self.h)

    def run_kernel_time(self):
        ms, k = C.c_double(), C.c_int64()
        self._chk(self.lib.grlcu_run_kernel_time(self.h, C.byref(ms), C.byref(k)))
        return ms.value, k.value

    def replay_step(self, inp, want_trace=True):
        """inp: dict of SoA numpy arrays named like grlcu_replay_in; returns dict of outputs."""
        n, S, OBS, T, A = self.n, self.S, self.cfg.obs_dims, self.cfg.tilings, self.cfg.n_actions
        HA = A
        if self.cfg.algo == ALGO_AC:
            A, HA = 4, 1       # see grlcu_replay_step in include/grlcu.h
        ri = ReplayIn()
        keep = []
        spec = dict(started=np.int32, start_state=np.float64, start_u=np.float64, start_r_explore=np.uint32,
                    start_r_tie=np.uint32, state=np.float64, action=np.float64, test=np.int32,
                    obs=np.float64, reward=np.float64, terminal=np.int32, u=np.float64,
                    r_explore=np.uint32, r_tie=np.uint32, force_action=np.int32)
        ptrty = {np.int32: _i32p, np.float64: _dp, np.uint32: _u32p}
        for k, dt in spec.items():
            v = inp.get(k)
            if v is None:
                continue
            arr = np.ascontiguousarray(v, dtype=dt)
            keep.append(arr)
            setattr(ri, k, _ptr(arr, ptrty[dt]))
        out = dict(next=np.empty((S, n)), obs=np.empty((OBS, n)), reward=np.empty(n), terminal=np.empty(n, np.int32),
                   hash=np.zeros((HA, T, n), np.uint32), index=np.zeros((HA, T, n), np.uint32), q=np.zeros((A, n)),
                   action_index=np.empty(n, np.int32), start_q=np.zeros((A, n)),
                   start_action_index=np.zeros(n, np.int32), q_prev=np.empty(n), target=np.empty(n),
                   delta=np.empty(n), swept=np.empty(n, np.int32), trace_len=np.empty(n, np.int32))
        if want_trace:
            out["trace_index"] = np.empty((self.trace_cap, T, n), np.uint32)
        ro = ReplayOut()
        for k, arr in out.items():
            setattr(ro, k, _ptr(arr, ptrty[arr.dtype.type]))
        self._chk(self.lib.grlcu_replay_step(self.h, C.byref(ri), C.byref(ro)))
        return out
