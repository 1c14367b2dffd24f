#!/usr/bin/env python
"""bench.py -- agent-steps/sec of the batched GRL learning loop (step + tilecode + SARSA(lambda) update).

One bench "step" = one pass of the hot path over one synthetic start-state batch: every agent runs
STEPS_PER_PASS (=100, one pendulum episode) agent-steps.  Workload at N=1: BASELINE.json configs[1],
pendulum swing-up SARSA(lambda) tile coding, 65,536 independent agents on one B200, per-agent table
memory=65,536 entries (32 GiB of tables: far larger than L2, so nothing is cache-warm between passes).

  value  : device-resident throughput (inputs already in HBM), CUDA events on the launching stream
  e2e    : the same through grlcu_run_host with HOST buffers (H2D start states + D2H results inside)
  roofline: algorithmic table bytes (640 + 16*V per agent-step, V measured on device) / run-kernel time
  cpu_baseline: the reference's own deployer built from its sources (oracle/_ref/grld_ref; Eigen, absent here, is a stand-in),
            one process per host core; the CPU oracle (port of the reference) beside it, and instead of it where the binary is absent

--impl reference times that CPU implementation alone, with the same JSON line layout.
"""
import argparse
import ctypes as C
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

STEPS_PER_PASS = 100
METRIC = "agent-steps/sec (step+tilecode+SARSA update)"

# BASELINE.json configs: [1] is the headline (default); the others are selectable for the result tables.
#   name: (config maker, agents per GPU, table entries per agent, table bytes formula, description)
WORKLOADS = {
    "pendulum": ("pendulum_config", 65536, 65536,
                 "pendulum swing-up SARSA(lambda) tile coding (cfg/pendulum/sarsa_tc.yaml)"),
    "cart_pole": ("cart_pole_config", 32768, 262144,
                  "cart-pole swing-up SARSA(lambda) tile coding (cfg/cart_pole/sarsa.yaml), 4-D state, 3 actions"),
    "cart_double_pole_ac": ("cart_double_pole_ac_config", 65536, 131072,
                            "cart-double-pole actor-critic, two tile-coded tables per agent (cfg/cart_double_pole/ac.yaml)"),
    "cart_double_pole": ("cart_double_pole_config", 65536, 131072,
                         "cart-double-pole balancing SARSA(lambda) tile coding (cfg/cart_double_pole/sarsa.yaml), 6-D state, "
                         "lambda 0.92: 51-entry trace"),
    "pendulum_ac": ("pendulum_ac_config", 65536, 65536,
                    "pendulum swing-up actor-critic, two tile-coded tables per agent (cfg/pendulum/ac_tc.yaml)"),
    "cart_pole_ac": ("cart_pole_ac_config", 32768, 262144,
                     "cart-pole swing-up actor-critic, two tile-coded tables per agent (cfg/cart_pole/ac_tc.yaml)"),
    "sweep": ("pendulum_config", 131072, 32768,
              "pendulum SARSA-TC alpha x lambda x epsilon sweep (4 x 8 x 4 grid), per-agent hyper-parameters"),
}
AC_WORKLOADS = ("cart_double_pole_ac", "pendulum_ac", "cart_pole_ac")
SWEEP_ALPHA = [0.05, 0.1, 0.2, 0.4]
SWEEP_LAMBDA = [0.0, 0.2, 0.4, 0.5, 0.65, 0.75, 0.85, 0.9]
SWEEP_EPSILON = [0.01, 0.02, 0.05, 0.1]


def make_config(mod, args):
    """The workload's config from `mod` (grl_b200 or the oracle harness: same maker names, same fields)."""
    maker = WORKLOADS[args.workload][0]
    kw = dict(memory=args.memory, test_interval=-1)
    if args.workload not in ("cart_double_pole_ac", "cart_double_pole"):       # (task/cart_double_pole/balancing has no randomization parameter)
        kw["randomization"] = 1.0
    if args.workload == "sweep":
        kw["lambda_"] = max(SWEEP_LAMBDA)          # trace ring sized for the largest lambda of the grid
    return getattr(mod, maker)(**kw)


def sweep_hyper(first_id, n):
    """Per-agent (alpha, lambda, epsilon) of the 4 x 8 x 4 grid, keyed by the global agent id.  lambda is constant
    over each block of 32 consecutive agents (a warp sweeps as long as its longest trace): block b = g // 32 runs
    lambda[b % 8]; alpha and epsilon cycle through the agents that share a lambda.  Every cell gets the same number
    of agents.  The same rule as experiment/online_learning's sweep_* parameters (grl_b200/host/grl_experiment.cpp)."""
    import numpy as np
    gid = np.arange(first_id, first_id + n)
    na, nl = len(SWEEP_ALPHA), len(SWEEP_LAMBDA)
    blk, r = gid // 32, gid % 32
    k = (blk // nl) * 32 + r
    a = np.array(SWEEP_ALPHA)[k % na]
    l = np.array(SWEEP_LAMBDA)[blk % nl]
    e = np.array(SWEEP_EPSILON)[(k // na) % len(SWEEP_EPSILON)]
    return a, l, e


def table_bytes_per_step(workload, V):
    """algorithmic table bytes per agent-step (8-byte entries, each touched entry once per read and
    once per write): SARSA 8*A*T + 8*T + 8*T + 16*V = 640 + 16 V; AC 8*T + 16*T + 2*8*T + 8*T + 16*V = 768 + 16 V."""
    return (768 if workload in AC_WORKLOADS else 640) + 16 * V


def committed_traffic(workload, agents, memory, agent_steps_per_launch):
    """DRAM traffic of one launch of the run kernel, from the committed `ncu --set full` capture of this exact workload
    (profiles/traffic.json, written by tools/update_traffic.py): (bytes per launch, capture file, sha of the kernel
    sources the capture was taken from, whether that is the current tree).  Per launch, like `roofline.achieved`: the
    capture's bytes per agent-step x the agent-steps of one launch of THIS run (equal to the capture's own
    dram_bytes_per_launch when the launches hold the same agents).  A cross-reference, not a measurement of the run
    that prints it: it is only turned into `traffic_frac` when the kernel sources are byte-identical."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get("%s:%d:%d" % (workload, agents, memory))
    except (OSError, ValueError) as exc:
        sys.stderr.write("bench: profiles/traffic.json not usable (%s)\n" % exc)
        return None, None, None, False
    if not t:
        return None, None, None, False
    tree = t.get("kernel_sources_sha")
    return (t["dram_bytes_per_agent_step"] * agent_steps_per_launch, t["capture"], tree,
            tree == kernel_sources_sha())


def kernel_sources_sha():
    """sha1 over the CUDA sources of libgrlcu.so: ties a committed ncu capture to the tree it was taken from."""
    import hashlib
    h = hashlib.sha1()
    d = os.path.join(ROOT, "grl_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------- CPU oracle timing (reference arm)
def _cpu_worker(wargs):
    args, agent_id, n_steps = wargs
    import _oracle as O
    L = O.lib()
    cfg = make_config(O, args)
    a = L.orc_create(C.byref(cfg), agent_id)
    if args.workload == "sweep":
        al, la, ep = sweep_hyper(agent_id, 1)
        L.orc_set_hyper(a, float(al[0]), float(la[0]), float(ep[0]))
    L.orc_run(a, 2000)                         # warm the table/caches
    t0 = time.perf_counter()
    done = L.orc_run(a, n_steps)
    dt = time.perf_counter() - t0
    L.orc_destroy(a)
    return done, dt


def cpu_oracle_rate(args, steps_per_proc, procs=None):
    """One oracle process per host core (the reference's grld is single-agent, single-thread);
    aggregate agent-steps / wall time of the slowest process."""
    import _oracle as O
    O.build()
    procs = procs or os.cpu_count() or 1
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(procs) as pool:
        res = pool.map(_cpu_worker, [(args, i, steps_per_proc) for i in range(procs)])
    wall = time.perf_counter() - t0
    total = sum(r[0] for r in res)
    slowest = max(r[1] for r in res)
    return total / slowest, procs, wall


GRLD_REF = os.path.join(ROOT, "oracle", "_ref", "grld_ref")


def _grld_batch(args, procs, trials, top, limit=300.0):
    """`procs` concurrent `grld_ref -s <1+i> cfg.yaml` processes, each in its own directory (one grld per host
    core, test_interval -1).  Returns (learning steps done by all, wall seconds from the first launch to the last exit)."""
    import _oracle as O
    runs = []
    for i in range(procs):
        d = os.path.join(top, "p%d_%d" % (i, trials))
        os.makedirs(d, exist_ok=True)
        cfg = make_config(O, args)
        if args.workload == "sweep":
            al, la, ep = sweep_hyper(i * 32, 1)                # block i: lambda cycles through the grid
            cfg.alpha, cfg.lambda_, cfg.epsilon = float(al[0]), float(la[0]), float(ep[0])
        with open(os.path.join(d, "cfg.yaml"), "w") as f:
            f.write(O.grld_yaml(cfg, trials, "curve"))
        runs.append(d)
    t0 = time.perf_counter()
    ps = [subprocess.Popen([GRLD_REF, "-s", str(1 + i), "cfg.yaml"], cwd=d, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
          for i, d in enumerate(runs)]
    try:
        for p in ps:
            if p.wait(timeout=max(limit - (time.perf_counter() - t0), 1.0)) != 0:
                raise RuntimeError("grld_ref exited with %d" % p.returncode)
    finally:
        for p in ps:
            if p.poll() is None:
                p.kill()
    wall = time.perf_counter() - t0
    steps = 0
    for d in runs:
        rows = [ln.split() for ln in open(os.path.join(d, "curve-0.txt")) if ln.strip()]
        steps += int(rows[-1][1])                              # column 2: learning steps so far (online_learning.cpp:252)
    return steps, wall


def grld_rate(args, seconds):
    """The reference's OWN deployer -- oracle/_ref/grld_ref: grld built from the reference's translation units with its own
    headers, flags and vendored yaml-cpp, Eigen being a stand-in -- one process per host core on the
    workload's configuration, about `seconds` of work.  None where the binary is absent or does not run."""
    if not os.path.exists(GRLD_REF):
        return None
    import tempfile
    try:
        procs = os.cpu_count() or 1
        with tempfile.TemporaryDirectory() as top:
            # two short runs separate the start-up (YAML, instantiation, table initialisation) from the per-step cost
            s0, w0 = _grld_batch(args, procs, 100, top, limit=60.0)
            s1, w1 = _grld_batch(args, procs, 600, top, limit=120.0)
            # (never trust a difference of two noisy wall times alone: bounded below by a quarter of the plain average)
            per_trial = max((w1 - w0) / 500.0, w1 / 600.0 / 4.0, 1e-6)
            trials = max(int(seconds / per_trial), 1000)
            steps, wall = _grld_batch(args, procs, trials, top, limit=30.0 + 20.0 * seconds)
        value = (steps - s1) / max(wall - w1, 1e-6)                # differential against the (warm) 600-trial run: start-up excluded
        return {"value": value, "unit": "agent-steps/s", "cores": procs, "kind": "reference",
                "sample": "%d grld_ref processes (one per core, -s 1..%d) x %d trials = %d agent-steps in %.1f s, rate taken as the "
                          "difference to a 600-trial run of the same processes (%d agent-steps in %.2f s) so that start-up does not "
                          "count; memory=%d, test_interval -1; grld built from the reference's own sources, Eigen replaced by a "
                          "stand-in with the same in-place storage" % (procs, procs, trials, steps, wall, s1, w1, args.memory)}
    except Exception as exc:
        sys.stderr.write("bench: grld_ref not usable (%s); the oracle port is timed instead\n" % exc)
        return None


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # the reference's own deployer where it was built (oracle/_ref/grld_ref), one process per core; else the oracle port
    samples, t0 = [], time.perf_counter()
    for k in range(args.warmup + args.steps):
        if k == args.warmup:
            t0 = time.perf_counter()
        r = grld_rate(args, 4.0)                                # ~4 s of CPU work per bench step (+ two short sizing runs)
        if r is None:
            samples = None
            break
        if k >= args.warmup:
            samples.append(r)
    ref_wall = time.perf_counter() - t0
    # the port, always: the other end of the bracket (no per-step heap traffic)
    rate0, procs, _ = cpu_oracle_rate(args, 50000)
    per_proc = max(int(rate0 / procs * 4.0), 10000)
    if samples:
        port_rate, procs, _ = cpu_oracle_rate(args, per_proc)
        wall = ref_wall
        value = sum(x["value"] for x in samples) / len(samples)
        baseline = dict(samples[-1], value=value,
                        port={"value": port_rate, "unit": "agent-steps/s", "cores": procs,
                              "sample": "%d oracle processes x %d agent-steps, memory=%d" % (procs, per_proc, args.memory)})
        note = ("the timed CPU implementation is the reference's own deployer built from its own sources (oracle/_ref/grld_ref; "
                "the one stand-in is Eigen, absent here), whose learning curves the oracle reproduces "
                "(tests/test_oracle_vs_grld_ref.py); cpu_baseline.port is the oracle port on the same cores")
    else:
        for _ in range(args.warmup):
            cpu_oracle_rate(args, per_proc // 4)
        rates, t0 = [], time.perf_counter()
        for _ in range(args.steps):
            r, procs, _ = cpu_oracle_rate(args, per_proc)
            rates.append(r)
        wall = time.perf_counter() - t0
        value = sum(rates) / len(rates)
        baseline = {"value": value, "unit": "agent-steps/s", "cores": procs, "kind": "port",
                    "sample": "%d oracle processes x %d agent-steps per bench step, memory=%d" % (procs, per_proc, args.memory)}
        note = ("oracle/_ref/grld_ref is not here (it is built where /root/reference exists): the timed CPU implementation is "
                "the oracle port, pinned to the reference's golden curve (tests/test_oracle_golden.py)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "agent-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * wall / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, max(args.gpus, 1)),
        "cpu_baseline": baseline,
        "e2e": {"value": value, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": note,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, n_gpus):
    tables = 2 if args.workload in AC_WORKLOADS else 1
    cfg = {"workload": "%s, %d independent agents per GPU" % (WORKLOADS[args.workload][3], args.agents),
           "name": args.workload, "agents_per_gpu": args.agents, "agents_total": args.agents * n_gpus,
           "memory_per_agent": args.memory, "tables_per_agent": tables,
           "agent_steps_per_pass": STEPS_PER_PASS, "tilings": 16, "trace": "replacing",
           "test_interval": -1, "cohort": args.cohort,
           "l2": "tables (agents*memory*8 B*%d = %.1f GiB per GPU) are far larger than L2; no flush needed"
                 % (tables, tables * args.agents * args.memory * 8 / 2**30)}
    if args.workload == "sweep":
        cfg["grid"] = {"alpha": SWEEP_ALPHA, "lambda": SWEEP_LAMBDA, "epsilon": SWEEP_EPSILON}
    return cfg


# ---------------------------------------------------------------- clocks sampler
class Clocks(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.rows.append([x.strip() for x in out.stdout.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]),
                "reasons": reasons, "samples": len(self.rows)}


# ---------------------------------------------------------------- our arm
def ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import grl_b200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the GRL B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cfg = make_config(grl_b200, args)
    stream = torch.cuda.current_stream().cuda_stream
    from grl_b200.dist import shard
    first_id, _ = shard(rank, world, args.agents)
    batch = grl_b200.Batch(cfg, args.agents, first_agent_id=first_id, device=local, stream=stream)
    n, S = args.agents, batch.S
    layout = batch.layout_info()
    bufs = batch.result_buffers()
    if args.workload == "sweep":
        batch.set_hyper(*sweep_hyper(first_id, n))

    # ---- device-resident throughput
    for _ in range(args.warmup):
        batch.run(STEPS_PER_PASS, args.cohort)
    batch.sync()
    batch.run_kernel_time()                    # drop warm-up events
    r0 = batch.results()
    launches0 = batch.launch_count()
    clocks = Clocks(local)
    clocks.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        batch.run(STEPS_PER_PASS, args.cohort)
    e1.record()
    barrier()
    dev_ms = e0.elapsed_time(e1)
    kern_ms, kern_launches = batch.run_kernel_time()
    launches = batch.launch_count() - launches0
    r1 = batch.results()
    steps_done = int((r1["steps"] - r0["steps"]).sum())
    swept = int((r1["swept"] - r0["swept"]).sum())
    writes = int((r1["table_writes"] - r0["table_writes"]).sum())

    # ---- end to end through the host-buffer call: every step copies ITS OWN start-state batch in from pinned host
    # memory and its per-agent results out into pinned host memory (the batches are generated before the clock starts).
    # An agent takes its row when it BEGINS its next episode; episodes in progress continue, so this leg runs the same
    # trajectory phase as the device-resident one (it does not restart every agent every step)
    rng = np.random.default_rng(1234 + rank)
    n_batches = max(args.steps, 1)
    starts = torch.zeros((n_batches, n, S), dtype=torch.float64).pin_memory()
    starts_np = starts.numpy()
    # Task::start of the workload's task on the host (pendulum.cpp:97-105, cart_pole.cpp:155-166,
    # cart_double_pole.cpp:250-262): synthetic start-state batches for the end-to-end leg
    if cfg.dynamics == 0:
        starts_np[:, :, 0] = np.pi + rng.uniform(0, 2 * np.pi, (n_batches, n))
    elif cfg.dynamics == 1:
        starts_np[:, :, 1] = np.pi + (0.1 * rng.uniform(0, 1, (n_batches, n)) - 0.05)
    else:
        starts_np[:, :, 1] = 0.01 * rng.uniform(0, 1, (n_batches, n)) - 0.005
    pinned = {k: torch.zeros(v.shape, dtype=getattr(torch, str(v.dtype))).pin_memory() for k, v in bufs.items()}
    pbufs = {k: v.numpy() for k, v in pinned.items()}
    # Same trajectory phase as the device-resident leg: throughput drifts as the agents learn (trace contents, re-visited
    # tiles), so the agents are reset (tables, traces, counters) and the leg repeats the schedule of the first one -- W
    # untimed warm-up passes, then the K timed passes -- through the host-buffer call.
    batch.reset()
    for _ in range(args.warmup):
        batch.run_host(starts_np[0], STEPS_PER_PASS, args.cohort, pbufs)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        batch.run_host(starts_np[k], STEPS_PER_PASS, args.cohort, pbufs)
        _ = float(pbufs["last_return"][0])
    barrier()
    e2e_s = time.perf_counter() - t0
    clocks.stop_flag = True
    clocks.join(timeout=2)
    h2d = n * S * 8
    d2h = n * (8 * 3 + 8 + 4 + 8 + 8)

    # ---- max over ranks; the only collective: final statistics gather over NCCL (outside the timed region; timed
    # here on the device so that the line shows what is outside the clock)
    from grl_b200.dist import finalize
    barrier()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    (dev_ms, e2e_s, kern_ms), (steps_all, swept_all, writes_all), ret = finalize(
        [dev_ms, e2e_s, kern_ms], [steps_done, swept, writes], r1["last_return"], "cuda")
    c1.record()
    torch.cuda.synchronize()
    collective_ms = c0.elapsed_time(c1)
    mean_return = float(ret.mean())

    if rank == 0:
        peak, peak_src = peaks()
        V = swept_all / max(steps_all, 1)
        B = table_bytes_per_step(args.workload, V)
        per_gpu_steps = steps_all / world
        achieved = per_gpu_steps * B / (kern_ms * 1e-3) / 1e9 if kern_ms > 0 else 0.0
        traffic, traffic_src, traffic_tree, traffic_current = committed_traffic(
            args.workload, args.agents, args.memory, per_gpu_steps / max(kern_launches, 1))
        # What actually bounds the path: every table access is a hash-random 8-byte word.  Ceilings measured on
        # this GPU type (profiles/random_access_ceilings.json); accesses counted on the device.
        random_access = None
        try:
            # the ceilings of THIS device, measured now (grlcu_probe_random_access: independent random 8-byte reads /
            # writes over an 8 GiB scratch buffer); the committed figures are the fallback
            try:
                r_s, w_s, _ = grl_b200.capi.probe_random_access(local, 1 << 33)
                ceil = {"read_G_per_s": r_s / 1e9, "write_G_per_s": w_s / 1e9,
                        "source": "measured on this device after the timed region (grlcu_probe_random_access, 8 GiB footprint)"}
            except Exception:
                with open(os.path.join(ROOT, "profiles", "random_access_ceilings.json")) as f:
                    ceil = json.load(f)
                ceil["source"] = "profiles/random_access_ceilings.json"
            ac = args.workload in AC_WORKLOADS
            cached = os.environ.get("GRLCU_MODE", "0") == "0"
            W = writes_all / max(steps_all, 1)
            R = (4 * 16 if ac else 3 * 16 + 16) + (0 if cached else V)      # cached path: the sweep runs on the dense cache
            if layout["tile_major"]:
                # one sector request per tiling for Q(s', .), one more per word that shares an earlier tile's slot, T for Q(p)
                R = 16 + 48.0 * layout["shared_words"] / max(layout["inbox_words"], 1) + 16
            floor_ns = R / ceil["read_G_per_s"] + W / ceil["write_G_per_s"]
            got_ns = kern_ms * 1e6 / max(per_gpu_steps, 1)
            random_access = {"kind": "model estimate, not a ceiling: the probe's addresses are uniformly random, the "
                                     "kernel's are clustered (a fraction above 1 means the clustering pays)",
                             "reads_per_agent_step": R, "writes_per_agent_step": W,
                             "read_ceiling_G_per_s": ceil["read_G_per_s"], "write_ceiling_G_per_s": ceil["write_G_per_s"],
                             "ns_per_agent_step_at_ceiling": floor_ns, "ns_per_agent_step": got_ns,
                             "frac": floor_ns / got_ns if got_ns > 0 else None,
                             "ceiling_source": ceil.get("source"),
                             "note": "time the measured counts of random 8-byte table reads and writes would take at the probed rates, over the "
                                     "kernel's measured time; the timed region leaves the trace value cache live (its write-back to the "
                                     "tables, k_cache_close, happens when the host next looks and is outside the clock)"}
        except Exception:
            pass
        value = steps_all / (dev_ms * 1e-3)
        e2e_val = args.agents * world * STEPS_PER_PASS * args.steps / e2e_s
        line = {
            "metric": METRIC, "value": value, "unit": "agent-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / max(args.steps, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world),
            "e2e": {"value": e2e_val, "unit": "agent-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            "clocks": clocks.summary(),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         # sector-granular figure: the DRAM bytes the hardware moved per launch (ncu capture)
                         # over this run's measured launch time, as a fraction of the same peak
                         "traffic_tree": traffic_tree, "traffic_is_of_this_tree": traffic_current,
                         "traffic_frac": (traffic / (kern_ms / max(kern_launches, 1) * 1e-3) / 1e9 / peak)
                                         if traffic and kern_ms > 0 and traffic_current else None,
                         "peak_source": peak_src, "random_access": random_access,
                         "bytes_per_agent_step": B, "V_mean": V,
                         "table_layout": dict(layout, note="tile-major / visit-ordered: the 3 action words of one tile share a 32-byte sector "
                                              "(DESIGN.md section 4); inbox_words = tile words ranked, shared_words = those sharing an earlier slot"),
                         "kernel": "grlcu::k_run_ac" if args.workload in AC_WORKLOADS else "grlcu::k_run",
                         "kernel_ms_per_launch": kern_ms / max(kern_launches, 1),
                         "kernel_ms_per_pass": kern_ms / max(args.steps, 1),
                         "kernel_launches": int(kern_launches)},
            "mean_last_return": mean_return,
            "collective_ms": collective_ms,
            "collective": "one all-gather of the per-agent returns + two all-reduces (timings MAX, counters SUM) after the "
                          "timed region (grl_b200/dist.py); nothing collective runs inside it",
        }
        if world == 1 and not args.no_cpu:
            rate0, procs, _ = cpu_oracle_rate(args, 50000)
            per_proc = max(int(rate0 / procs * 12.0), 10000)
            rate, procs, wall = cpu_oracle_rate(args, per_proc)
            port = {"value": rate, "unit": "agent-steps/s", "cores": procs, "kind": "port",
                    "sample": "%d oracle processes (one per core) x %d agent-steps, memory=%d, %.1f s"
                              % (procs, per_proc, args.memory, wall)}
            ref = grld_rate(args, 10.0)                         # the reference's own deployer, where it was built
            line["cpu_baseline"] = dict(ref, port=port) if ref else port
        print(json.dumps(line), flush=True)
    batch.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="pendulum", choices=sorted(WORKLOADS))
    ap.add_argument("--agents", type=int, default=None, help="agents per GPU (default: the workload's)")
    ap.add_argument("--memory", type=int, default=None, help="table entries per agent (default: the workload's)")
    ap.add_argument("--cohort", type=int, default=0, help="agents per L2-resident cohort (0 = all at once)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.agents is None:
        args.agents = WORKLOADS[args.workload][1]
    if args.memory is None:
        args.memory = WORKLOADS[args.workload][2]
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours(args)


if __name__ == "__main__":
    main()
