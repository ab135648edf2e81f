#!/usr/bin/env python
"""bench.py — agent-steps/s of the batched MARL-Coverage env step on B200 (contract: task brief).

One "step" = one pass of the hot path (DecGridRL.step for every env of this GPU's shard, with
auto-reset and episode statistics) over one batch of synthetic actions.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c2|c3|c4|super] [--impl reference]

Default workload "c5" = BASELINE.json configs[4]: weak scaling, 131 072 envs per GPU x 4 agents on
bg2_100x100, per-env settings of configs[1] (square sensor range 2, egoradius 5, obs [4,3,11,11] f32).
State + observations are ~2.9 GB per GPU, far larger than the 126 MB L2, so no L2 flush is needed.
The default JSON line also carries `workloads`: configs[1] at its literal size (4096 envs), configs[2]
(lidar), configs[3] (1073^2 maps) and the SuperGridRL variant, each as us per step of the step kernel and
fraction of the HBM roofline (small batches: L2 flushed between individually timed steps).

`--impl reference` times the CPU implementation of the same path (the numpy oracle port of the
pure-Python reference — the reference itself cannot travel to the GPU box; profiles/cpu_reference_vs_port.json
holds the measured speed relation of the two) on all host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# steps per statistics all-reduce at N > 1 (every step's vector is reduced, STATS_BLOCK of them per call)
STATS_BLOCK = max(1, int(os.environ.get("MCOV_STATS_BLOCK", "8")))
METRIC = "agent-steps/sec (batched env step, 1/2/4/8 B200) and % of HBM roofline"
UNIT = "agent-steps/s"

DEC_BASE = dict(maxsteps=1000, collision_penalty=5, done_thresh=1, done_incr=0, terminal_reward=30,
                dist_reward=0, train_maxsteps=1000, test_maxsteps=1000, egoradius=5, mini_map_rad=0,
                comm_radius=0, allow_comm=0, map_sharing=0, single_square_tool=0, dijkstra_input=0)
C2_CFG = dict(DEC_BASE, numrobot=4, sensor_type="square_sensor", sensor_config={"range": 2})

WORKLOADS = {
    # name: (env kind, grid set, envs per GPU, env_config, description)
    "c5": ("dec", "bg2_100x100", 131072, C2_CFG,
           "configs[4] weak scaling: 131072 envs/GPU x 4 agents, DecGridRL bg2_100x100, square sensor "
           "r=2, egoradius 5 (per-env settings of configs[1])"),
    "c2": ("dec", "bg2_100x100", 4096, C2_CFG,
           "configs[1]: 4096 envs x 4 agents, DecGridRL bg2_100x100, square sensor r=2, egoradius 5"),
    "c3": ("dec", "bg2_256x256", 16384,
           dict(DEC_BASE, numrobot=8, sensor_type="lidar", sensor_config={"range": 10, "num_lasers": 21}),
           "configs[2]: 16384 envs x 8 agents, DecGridRL bg2_256x256, lidar 21 lasers range 10"),
    "c4": ("dec", "bg2_1073x1073", 256,
           dict(DEC_BASE, numrobot=16, sensor_type="square_sensor", sensor_config={"range": 2}),
           "configs[3]: 256 envs x 16 agents, DecGridRL bg2_1073x1073, square sensor r=2"),
    "super": ("super", "bg2_100x100", 8192,
              dict(numrobot=4, train_maxsteps=1000, test_maxsteps=1000, collision_penalty=5,
                   senseradius=2, free_penalty=0.2, done_thresh=1, done_incr=0, terminal_reward=30,
                   dist_reward=0, use_scanning=0),
              "C1 settings batched: 8192 envs x 4 agents, SuperGridRL bg2_100x100, senseradius 2, "
              "full [7,100,100] f32 state"),
    # gated options of DecGridRL at configs[1] size: extra launches per step
    "c2_share": ("dec", "bg2_100x100", 4096, dict(C2_CFG, map_sharing=1, comm_radius=12, allow_comm=1),
                 "configs[1] size with map_sharing=1, comm_radius=12 (shareMaps launch before the step kernel)"),
    "c2_dist": ("dec", "bg2_100x100", 4096, dict(C2_CFG, dist_reward=1),
                "configs[1] size with dist_reward=1 (move / distance maps / sense+observe: the split step)"),
    "c2_dijkstra": ("dec", "bg2_100x100", 4096, dict(C2_CFG, dijkstra_input=1),
                    "configs[1] size with dijkstra_input=1 (bit-plane BFS path layer after the step kernel)"),
}


def alg_bytes_per_env_step(kind, cfg, W, L):
    """SURVEY.md §8(d) algorithmic bytes per env-step (the roofline numerator)."""
    N = cfg["numrobot"]
    if kind == "super":
        return (N + 3) * W * L * 4 + N * 1 + 5 + N * 8
    e = cfg["egoradius"]
    ow = 2 * e + 1
    C = 3
    R = cfg["sensor_config"]["range"]
    S = 2 * R + 1
    b_obs = N * C * ow * ow * 4
    per_robot_rows = 3 * ow + 2 * S + S + 2 * S
    if cfg["sensor_type"] == "lidar":
        per_robot_rows += 2 * S
    b_state = N * 8 * per_robot_rows + N * 16
    return b_obs + N * 1 + 5 + N * 8 + b_state


def plane_words(W, L):
    return ((L + 31) // 32) * ((W + 3) // 4 * 4)


def obs_shape_of(kind, cfg, E, W, L):
    N = cfg["numrobot"]
    if kind == "super":
        return [E, 4 if cfg.get("use_scanning") else N + 3, W, L]
    ow = 2 * cfg["egoradius"] + 1
    C = (5 if cfg.get("mini_map_rad", 0) > 0 else 3) + int(bool(cfg.get("dist_reward"))) + \
        int(bool(cfg.get("dijkstra_input")))
    return [E, N, C, ow, ow]


def working_set_bytes(kind, cfg, E, W, L):
    """State planes + observation block of one GPU's shard (W, L in env coordinates)."""
    N = cfg["numrobot"]
    obs = 4
    for d in obs_shape_of(kind, cfg, E, W, L):
        obs *= d
    planes = (2 if kind == "super" else 1 + 2 * N) * plane_words(W, L) * 4
    return obs + E * planes


L2_BYTES = 126 * 1024 * 1024


def grid_set_info(name):
    """(number of maps, W, L) of a committed grid set without unpacking it."""
    import numpy as np
    z = np.load(os.path.join(ROOT, "data", "grids", name + ".npz"))
    W, L = (int(v) for v in z["shape"])
    return int(z["bits"].shape[0]), W, L


def config_dict(workload, E, world):
    """The `config` object of the JSON line — a function of the workload table, the shard size and the GPU
    count only, so the GPU arm and the reference arm print the same one.  Returns (config, l2_flushed)."""
    kind, grid_name, _, cfg, desc = WORKLOADS[workload]
    count, W, L = grid_set_info(grid_name)
    b = 2 if kind == "dec" else 0
    ws = working_set_bytes(kind, cfg, E, W + b, L + b)
    # what one step touches (windows + observation block), not what the batch owns: 256 envs on 1073^2 maps own
    # 1.2 GB of planes but a step reads and writes 8 MB of them, which would stay in the 126 MB L2
    touched = E * alg_bytes_per_env_step(kind, cfg, W + b, L + b)
    flushed = touched < 2 * L2_BYTES
    return {"workload": desc, "envs_per_gpu": E, "agents": cfg["numrobot"],
            "maps": "%s first %d sorted" % (grid_name, count // 2),
            "obs_shape": obs_shape_of(kind, cfg, E, W + b, L + b), "auto_reset": True,
            "truncate_steps": cfg["train_maxsteps"],
            "l2": ("flushed between timed steps (256 MiB write)" if flushed else
                   "working set %.2f GB per GPU (%.2f GB touched per step) >> 126 MB L2, no flush"
                   % (ws / 1e9, touched / 1e9)),
            "stats_allreduce": ("NCCL SUM of every step's statistics vector on a side stream, %d steps per "
                                "collective" % STATS_BLOCK) if world > 1 else "none (1 GPU)"}, flushed


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def port_vs_reference():
    """Measured speed of the oracle port relative to the unmodified reference env (same container, same
    settings; tools/cpu_reference_vs_port.py), or None when the record is missing."""
    try:
        with open(os.path.join(ROOT, "profiles", "cpu_reference_vs_port.json")) as f:
            d = json.load(f)
        return {"port_over_reference": d["summary"]["port_over_reference"],
                "source": "profiles/cpu_reference_vs_port.json"}
    except Exception:
        return None


# ---------------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="mcov_clocks_", suffix=".csv")
            os.close(fd)
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
        except Exception:
            pass
        try:
            self.f.close()
            rows = [r.strip().split(",") for r in open(self.path) if r.strip()]
            os.unlink(self.path)
        except Exception:
            return out
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if r[5 + k].strip().lower().startswith("active"):
                    reasons.add(nm)
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------- CPU (oracle) arm
_CPU = {}


def _cpu_init(kind, grid_name, cfg, seed_base):
    """Pool initializer: one env object per worker process, kept across samples."""
    import contextlib
    import io
    import numpy as np
    from marl_coverage_b200.maps import load_grid_set
    from oracle import coverage_oracle as oc
    grids = load_grid_set(grid_name)
    train = grids[:len(grids) // 2]
    seed = seed_base + os.getpid() % 100003
    np.random.seed(seed % (2 ** 31))
    with contextlib.redirect_stdout(io.StringIO()):
        env = oc.OracleDecGridRL(train, cfg) if kind == "dec" else oc.OracleSuperGridRL(train, cfg)
    _CPU.update(env=env, rng=np.random.default_rng(seed), N=cfg["numrobot"], maxsteps=cfg["train_maxsteps"])


def _cpu_sample(seconds):
    """Steps this worker's env for `seconds`; returns (env steps, elapsed)."""
    import contextlib
    import io
    env, rng, N, maxsteps = _CPU["env"], _CPU["rng"], _CPU["N"], _CPU["maxsteps"]
    steps = 0
    with contextlib.redirect_stdout(io.StringIO()):
        t0 = time.perf_counter()
        while True:
            a = int(rng.integers(0, 4 ** N)) if N <= 31 else rng.integers(0, 4, size=N)
            _, _, done = env.step(a)
            if done or env._currstep >= maxsteps:
                env.reset(False, None)
            steps += 1
            now = time.perf_counter()
            if now - t0 >= seconds:
                return steps, now - t0


class CpuArm:
    """Oracle port of the reference env on `procs` processes (one env object each, random base-4 actions,
    reset on done / train_maxsteps — BASELINE.md §3).  The pool and the env objects live across samples."""

    def __init__(self, kind, grid_name, cfg, procs=None):
        import multiprocessing as mp
        self.procs = procs or len(os.sched_getaffinity(0))
        self.N = cfg["numrobot"]
        self.pool = mp.get_context("spawn").Pool(self.procs, initializer=_cpu_init,
                                                 initargs=(kind, grid_name, cfg, 1000))

    def sample(self, seconds):
        """One bounded sample on every core.  Returns (agent-steps/s summed over workers, wall seconds)."""
        t0 = time.perf_counter()
        res = self.pool.map(_cpu_sample, [seconds] * self.procs, chunksize=1)
        wall = time.perf_counter() - t0
        return sum(s / t for s, t in res) * self.N, wall

    def close(self):
        self.pool.close()
        self.pool.join()


def run_reference(args):
    kind, grid_name, E, cfg, desc = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0
    if args.envs_per_gpu:
        E = args.envs_per_gpu
    # a "step" of this arm is one bounded sample: `sample_s` seconds of stepping on every host core; W untimed
    # samples, then exactly K timed ones, sized so that the whole run stays near two minutes
    steps, warm = max(1, args.steps), max(0, args.warmup)
    sample_s = args.sample_seconds or min(5.0, max(0.25, 110.0 / (steps + warm)))
    arm = CpuArm(kind, grid_name, cfg)
    for _ in range(warm):
        arm.sample(sample_s)
    rates, walls = [], []
    for _ in range(steps):
        r, w = arm.sample(sample_s)
        rates.append(r)
        walls.append(w)
    arm.close()
    value = sum(r * w for r, w in zip(rates, walls)) / sum(walls)
    config, _ = config_dict(args.workload, E, world)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": 1e3 * sum(walls) / len(walls),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config,
        "note": "CPU arm: a step of this arm = one bounded sample (%.2f s of stepping on every host core); numpy "
                "oracle port of the pure-Python reference env (the reference needs /root/reference, absent on the "
                "GPU box); one env object per process kept across samples, random base-4 actions, reset on done / "
                "train_maxsteps" % sample_s,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.procs, "kind": "port",
                         "sample": "%d x %.2f s of stepping on %d processes" % (steps, sample_s, arm.procs),
                         "vs_unmodified_reference": port_vs_reference()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------- GPU arm
def make_env(kind, grid_name, cfg, E, device, env_id_base, auto_reset=True):
    from marl_coverage_b200 import VecDecGridRL, VecSuperGridRL, load_grid_set
    grids = load_grid_set(grid_name)
    train = grids[:len(grids) // 2]                       # gridload half split (gridmaker.py:96-102)
    cls = VecDecGridRL if kind == "dec" else VecSuperGridRL
    env = cls(train, cfg, num_envs=E, device=device, auto_reset=auto_reset,
              truncate_steps=cfg["train_maxsteps"] if auto_reset else 0, env_id_base=env_id_base, seed=0)
    return env, len(train)


def gen_actions(torch, device, T, E, N, seed):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return torch.randint(0, 4, (T, E, N), dtype=torch.uint8, device=device, generator=g)


def median(xs):
    ys = sorted(xs)
    return ys[len(ys) // 2]


def timed_steps(torch, dist, env, acts, warmup, steps, world, flush=None, block=None):
    """W untimed + K timed steps with the statistics reduction in the loop.

    Returns (ms over the K steps by one CUDA-event pair on the launching stream, per-step ms list, totals,
    warm-up steps run).  Every code path of the timed region has run before the first event: the reducer is
    pre-warmed (one reduction of each block incl. the collective and torch's sum / add_ kernels,
    sharding.StatsReducer.warm) and the warm-up covers both statistics blocks.  With `flush` (a big tensor)
    the L2 is overwritten between steps and the total is the sum of the individually timed steps."""
    from marl_coverage_b200.sharding import StatsReducer
    dev = env.device
    block = block or STATS_BLOCK
    red = StatsReducer(dev, world=world, block=block)
    main = torch.cuda.current_stream(dev)
    warmup = max(warmup, 2 * block + 1)
    T = acts.shape[0]

    def one(k):
        env.stats = red.buffer()
        env.step(acts[k % T], want_stats=True)
        red.submit()

    for k in range(warmup):
        one(k)
    if world > 1:
        main.wait_stream(red.side)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(dev)
    if flush is None:
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        evs[0].record(main)
        for k in range(steps):
            one(warmup + k)
            if k + 1 < steps:
                evs[k + 1].record(main)
        red.flush()                                      # the timed steps' statistics are reduced inside the region
        if world > 1:
            main.wait_stream(red.side)
        evs[steps].record(main)
        torch.cuda.synchronize(dev)
        ms = evs[0].elapsed_time(evs[steps])
        per = [evs[k].elapsed_time(evs[k + 1]) for k in range(steps)]
    else:
        pairs = []
        for k in range(steps):
            flush.add_(1)                                # > L2-sized read+write
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(main)
            one(warmup + k)
            b.record(main)
            pairs.append((a, b))
        torch.cuda.synchronize(dev)
        per = [a.elapsed_time(b) for a, b in pairs]
        ms = sum(per)
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize(dev)
    red.finish()
    return ms, per, red.total, warmup


def kernel_only(torch, workload, device, rank, steps, flush_buf=None, E=None):
    """The dominant kernel alone: the step of a batch without auto-reset and without statistics (one launch
    per call for the plain configs), each call timed with its own CUDA-event pair on the launching stream."""
    kind, grid_name, E0, cfg, desc = WORKLOADS[workload]
    E = E or E0
    N = cfg["numrobot"]
    env, n_maps = make_env(kind, grid_name, cfg, E, device, rank * E, auto_reset=False)
    acts = gen_actions(torch, device, 32, E, N, 77 + rank)
    env.reset(map_idx=(torch.arange(E, device=device, dtype=torch.int32) + rank * E) % n_maps)
    _, flushed = config_dict(workload, E, 1)
    flush = flush_buf if flushed else None
    for k in range(8):
        env.step(acts[k % 32])
    torch.cuda.synchronize(device)
    main = torch.cuda.current_stream(device)
    pairs = []
    l0 = env.launch_count
    for k in range(steps):
        if flush is not None:
            flush.add_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(main)
        env.step(acts[(8 + k) % 32])
        b.record(main)
        pairs.append((a, b))
    torch.cuda.synchronize(device)
    launches = (env.launch_count - l0) / float(steps)
    per = [a.elapsed_time(b) for a, b in pairs]
    avg, med = sum(per) / len(per), median(per)
    peak, peak_src = measured_peak()
    balg = alg_bytes_per_env_step(kind, cfg, env.W, env.L)
    achieved = balg * E / (avg * 1e-3) / 1e9
    out = {"bound": "hbm", "kernel": "dec_env_kernel" if kind == "dec" else "super_env_kernel",
           "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
           "traffic": load_traffic(workload), "alg_bytes_per_agent_step": balg / N, "kernel_ms": avg,
           "kernel_ms_median": med, "launches_per_call": launches, "peak_source": peak_src,
           "l2": "flushed between launches" if flush is not None else "working set >> L2"}
    del env
    return out, E, N


def workload_summary(torch, workload, device, flush_buf, steps=60):
    """Short record of one of the other BASELINE configs for the default JSON line."""
    roof, E, N = kernel_only(torch, workload, device, 0, steps, flush_buf)
    med = roof["kernel_ms_median"]
    return {"workload": WORKLOADS[workload][4], "envs": E, "agents": N, "us_per_step": 1e3 * med,
            "value": E * N / (med * 1e-3), "unit": UNIT,
            "frac_of_hbm_peak": roof["frac"] * roof["kernel_ms"] / med,
            "launches_per_step": roof["launches_per_call"], "l2": roof["l2"]}


def run_gpu(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device; there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()

    kind, grid_name, E, cfg, desc = WORKLOADS[args.workload]
    if args.envs_per_gpu:
        E = args.envs_per_gpu
    N = cfg["numrobot"]
    K = args.steps
    config, flushed = config_dict(args.workload, E, world)

    env, n_maps = make_env(kind, grid_name, cfg, E, device, env_id_base=rank * E)
    T = 64
    acts = gen_actions(torch, device, T, E, N, 1234 + rank)
    map_idx = (torch.arange(E, device=device, dtype=torch.int32) + rank * E) % n_maps
    env.reset(map_idx=map_idx)                           # start poses: device Philox, seed 0
    torch.cuda.synchronize(device)

    flush_buf = torch.zeros(256 * 1024 * 1024 // 4, dtype=torch.int32, device=device)
    flush = flush_buf if flushed else None

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()

    # a second env batch of the same shape carries the load around the timed steps, so the measured batch
    # keeps its episode phase (all envs truncate together at multiples of train_maxsteps: a time-based soak
    # on the measured batch would move that mass-reset step in or out of the timed region at random)
    env_soak, _ = make_env(kind, grid_name, cfg, E, device, env_id_base=rank * E)
    env_soak.reset(map_idx=map_idx)

    def soak(seconds):
        """Untimed stepping load so that the 100 ms clock sampler sees the GPU in the state the timed steps
        run in (the timed region itself can be shorter than one sampling period)."""
        t_end = time.perf_counter() + seconds
        k = 0
        while time.perf_counter() < t_end:
            for _ in range(50):
                env_soak.step(acts[k % T])
                k += 1
            torch.cuda.synchronize(device)

    soak(0.6)
    launches0 = env.launch_count
    ms, per, total_stats, Wm = timed_steps(torch, dist, env, acts, max(3, args.warmup), K, world, flush=flush)
    launches = env.launch_count - launches0
    launches_timed = int(round(launches * K / float(K + Wm)))
    soak(0.4)
    clk = clocks.stop() if rank == 0 else None
    del env_soak
    if clk is not None:
        clk["window"] = "timed steps bracketed by 0.6 s + 0.4 s of the same stepping load (100 ms sampling period)"

    t = torch.tensor([ms, median(per)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, med_max = float(t[0].item()), float(t[1].item())
    value = world * E * N * K / (ms_max * 1e-3)

    # ---- K = 1 (a collective after every step) against K = STATS_BLOCK, same loop, same envs ------------
    block_ab = None
    if world > 1:
        Kb = min(K, 100)
        ab = {}
        for blk in sorted({1, STATS_BLOCK}):
            m, _, _, _ = timed_steps(torch, dist, env, acts, 3, Kb, world, flush=flush, block=blk)
            tt = torch.tensor([m], dtype=torch.float64, device=device)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ab["block_%d_ms_per_step" % blk] = float(tt.item()) / Kb
        block_ab = dict(ab, steps=Kb, note="same stepping loop, statistics all-reduced every `block` steps "
                                           "(block 1 = one NCCL collective per step)")

    # ---- dominant kernel alone (every rank runs it so ranks stay in lock step; rank 0 reports) ----------
    roof, _, _ = kernel_only(torch, args.workload, device, rank, min(K, 200), flush_buf, E=E)

    # ---- e2e: host buffers through the public API.  Every step: pinned host actions -> H2D, step,
    # reward / done / statistics -> D2H.  Double-buffered on two copy streams; the host waits for (= consumes)
    # the results of step t-1 while step t runs, as a rollout loop whose policy lives on the GPU does.  At
    # N > 1 the statistics go through the NCCL all-reduce first (StatsReducer, STATS_BLOCK steps per
    # collective on its side stream) and the host receives the GLOBAL rows of each block ------------------
    from marl_coverage_b200.sharding import StatsReducer
    main = torch.cuda.current_stream(device)
    acts_host = acts.cpu().pin_memory()
    act_dev = [torch.empty((E, N), dtype=torch.uint8, device=device) for _ in range(2)]
    res = [env.make_result_buffers() for _ in range(2)]
    host = [{"reward": torch.empty((E,), dtype=torch.float32).pin_memory(),
             "done": torch.empty((E,), dtype=torch.uint8).pin_memory(),
             "stats": torch.empty((8,), dtype=torch.float64).pin_memory()} for _ in range(2)]
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_step = [torch.cuda.Event() for _ in range(2)]
    ev_out = [torch.cuda.Event() for _ in range(2)]
    for ev in ev_step + ev_out:
        ev.record(main)
    Ke = min(K, 200)
    We = 2 * STATS_BLOCK + 2
    nblk = (Ke + We) // STATS_BLOCK + 4
    host_blocks = torch.zeros((nblk, STATS_BLOCK, 8), dtype=torch.float64).pin_memory()
    blk_ctr = [0]

    def to_host(view, first_step):
        host_blocks[blk_ctr[0] % nblk, :view.shape[0]].copy_(view, non_blocking=True)
        blk_ctr[0] += 1

    red = StatsReducer(device, world=world, block=STATS_BLOCK, on_reduced=to_host) if world > 1 else None

    def e2e_step(k):
        b = k & 1
        with torch.cuda.stream(s_in):
            s_in.wait_event(ev_step[b])                  # step k-2 has consumed act_dev[b]
            act_dev[b].copy_(acts_host[k % T], non_blocking=True)
            ev_in[b].record(s_in)
        main.wait_event(ev_in[b])
        main.wait_event(ev_out[b])                       # step k-2's results have left res[b]
        if red is not None:
            res[b]["stats"] = red.buffer()
        env.step(act_dev[b], want_stats=True, results=res[b])
        if red is not None:
            red.submit()
        ev_step[b].record(main)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_step[b])
            host[b]["reward"].copy_(res[b]["reward"], non_blocking=True)
            host[b]["done"].copy_(res[b]["done"], non_blocking=True)
            if red is None:
                host[b]["stats"].copy_(res[b]["stats"], non_blocking=True)
            ev_out[b].record(s_out)

    for k in range(We):
        e2e_step(k)
    if red is not None:
        red.flush()
        main.wait_stream(red.side)
    torch.cuda.synchronize(device)
    blk_ctr[0] = 0
    host_blocks.zero_()
    if world > 1:
        dist.barrier()
    host_sum = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    for k in range(Ke):
        e2e_step(We + k)
        if k >= 1:
            ev_out[(We + k - 1) & 1].synchronize()       # the caller reads step k-1's results here
            if red is None:
                host_sum += float(host[(We + k - 1) & 1]["stats"][6])
    ev_out[(We + Ke - 1) & 1].synchronize()
    if red is None:
        host_sum += float(host[(We + Ke - 1) & 1]["stats"][6])
    else:
        red.flush()
        main.wait_stream(red.side)
    main.wait_stream(s_out)
    e1.record(main)
    torch.cuda.synchronize(device)
    if red is not None:
        host_sum = float(host_blocks[:, :, 6].sum())
    if not os.environ.get("MCOV_DEBUG"):
        assert host_sum == float(Ke) * world * E * N, \
            "e2e: host saw %r agent-steps in the statistics, expected %r" % (host_sum, float(Ke) * world * E * N)
    ms_pipe = e0.elapsed_time(e1)
    if red is not None:
        red.finish()
    # the plain form of the same loop (one stream, host waits for every step's results): cheaper per call,
    # so it wins on small batches where a step is shorter than the extra stream / event calls
    ms_sync = float("inf")
    if world == 1:
        for k in range(3):
            act_dev[0].copy_(acts_host[k % T], non_blocking=True)
            env.step(act_dev[0], want_stats=True)
        torch.cuda.synchronize(device)
        e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e2.record(main)
        for k in range(Ke):
            act_dev[0].copy_(acts_host[k % T], non_blocking=True)
            _, r, d = env.step(act_dev[0], want_stats=True)
            host[0]["reward"].copy_(r, non_blocking=True)
            host[0]["done"].copy_(d, non_blocking=True)
            host[0]["stats"].copy_(env.stats, non_blocking=True)
            main.synchronize()
        e3.record(main)
        torch.cuda.synchronize(device)
        ms_sync = e2.elapsed_time(e3)
    te = torch.tensor([ms_pipe, ms_sync], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    ms_pipe, ms_sync = float(te[0].item()), float(te[1].item())
    e2e_ms = min(ms_pipe, ms_sync) / Ke
    e2e = {"value": world * E * N / (e2e_ms * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": E * N, "d2h_bytes_per_step": E * 4 + E + 64,
           "mode": "double-buffered" if ms_pipe <= ms_sync else "synchronous", "ms_per_step": e2e_ms,
           "ms_per_step_double_buffered": ms_pipe / Ke,
           "ms_per_step_synchronous": (ms_sync / Ke) if world == 1 else None,
           "stats_allreduce_in_loop": world > 1,
           "note": "observations stay in HBM for the policy; per step: actions H2D from pinned memory, "
                   "reward/done/stats D2H.  double-buffered = copies on two copy streams, the host reads "
                   "step t-1's results while step t runs; synchronous = one stream, host waits every step.  "
                   "N > 1: the statistics rows reach the host after the NCCL all-reduce of their block"}

    # ---- what the design avoids: the same loop with the observation block copied to the host as well -----
    obs_to_host = None
    if rank == 0 and world == 1 and not args.no_extras:
        try:
            obs_host = torch.empty(env.obs.shape, dtype=torch.float32).pin_memory()
            Ko = 6
            for k in range(2):
                env.step(act_dev[0], want_stats=True)
                obs_host.copy_(env.obs, non_blocking=True)
            torch.cuda.synchronize(device)
            e4, e5 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e4.record(main)
            for k in range(Ko):
                act_dev[0].copy_(acts_host[k % T], non_blocking=True)
                o, r, d = env.step(act_dev[0], want_stats=True)
                obs_host.copy_(o, non_blocking=True)
                host[0]["reward"].copy_(r, non_blocking=True)
                host[0]["done"].copy_(d, non_blocking=True)
                main.synchronize()
            e5.record(main)
            torch.cuda.synchronize(device)
            mo = e4.elapsed_time(e5) / Ko
            obs_to_host = {"value": E * N / (mo * 1e-3), "unit": UNIT, "ms_per_step": mo, "steps": Ko,
                           "d2h_bytes_per_step": int(env.obs.numel()) * 4 + E * 5,
                           "note": "secondary: the reference API's host-array return emulated by copying the "
                                   "whole float32 observation block D2H every step (PCIe-bound)"}
            del obs_host
        except Exception as ex:                              # the extra must never cost the bench line
            obs_to_host = {"error": repr(ex)[:200]}
    e2e["obs_to_host"] = obs_to_host

    others = None
    if rank == 0 and world == 1 and args.workload == "c5" and not args.no_extras:
        others = {}
        del env, res, act_dev
        torch.cuda.empty_cache()
        for wl in ("c2", "c3", "c4", "super", "c2_share", "c2_dist", "c2_dijkstra"):
            try:
                others[wl] = workload_summary(torch, wl, device, flush_buf)
            except Exception as ex:
                others[wl] = {"error": repr(ex)[:200]}
        for wl, replicas in (("c2", 6), ("c4", 24)):
            try:
                others[wl]["cuda_graph"] = run_graph_round_robin(torch, device, wl, replicas)
            except Exception as ex:
                others[wl]["cuda_graph"] = {"error": repr(ex)[:200]}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        arm = CpuArm(kind, grid_name, cfg)
        arm.sample(1.0)
        rate, _ = arm.sample(args.cpu_seconds)
        arm.close()
        cpu = {"value": rate, "unit": UNIT, "cores": arm.procs, "kind": "port",
               "sample": "numpy oracle port of the reference env, %d processes x %.0f s of stepping "
                         "(one env each, random base-4 actions, reset on done/train_maxsteps)"
                         % (arm.procs, args.cpu_seconds),
               "vs_unmodified_reference": port_vs_reference()}

    if rank == 0:
        st = total_stats.cpu().numpy()
        # value, its per-step median, e2e and the kernel alone all time the same step: they must agree
        rel_e2e = abs(ms_max / K - e2e_ms) / e2e_ms
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": ms_max / K, "ms_per_step_median": med_max, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 bit-planes (obs f32, reward f64)", "data": "synthetic",
            "config": config, "roofline": roof, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": launches_timed, "clocks": clk,
            "consistency": {"value_vs_e2e_rel_diff": rel_e2e,
                            "value_vs_kernel_ratio": (ms_max / K) / roof["kernel_ms"], "ok": bool(rel_e2e < 0.10),
                            "note": None if rel_e2e < 0.10 else
                            "value and e2e disagree by more than 10 %: look for a one-off cost inside a timed region"},
            "episode_stats": {"episodes": st[0], "mean_return": (st[1] / st[0]) if st[0] else None,
                              "mean_length": (st[2] / st[0]) if st[0] else None,
                              "agent_steps": st[6]},
        }
        if block_ab:
            line["stats_block_ab"] = block_ab
        if others:
            line["workloads"] = others
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_graph_round_robin(torch, device, workload, replicas, rounds=10, replays=5):
    """A small-batch step with launch overhead and cold instruction / constant caches taken out: `replicas`
    independent env batches stepped round-robin inside one CUDA graph (the C ABI only enqueues on the caller's
    stream, so a rollout captures as is).  The bytes the batches touch per round exceed the 126 MB L2, so every
    step finds its state cold without an explicit flush — while the kernel's code stays cached, as it does in a
    real rollout (the individually timed, L2-flushed figure next to it pays a cold start for both)."""
    kind, grid_name, E, cfg, _ = WORKLOADS[workload]
    N = cfg["numrobot"]
    envs = []
    n_maps = 1
    for r in range(replicas):
        env, n_maps = make_env(kind, grid_name, cfg, E, device, r * E)
        env.reset(map_idx=(torch.arange(E, device=device, dtype=torch.int32) + r) % n_maps)
        envs.append(env)
    acts = gen_actions(torch, device, rounds, replicas * E, N, 321).view(rounds, replicas, E, N)
    side = torch.cuda.Stream(device)
    side.wait_stream(torch.cuda.current_stream(device))
    with torch.cuda.stream(side):
        for r, env in enumerate(envs):
            env.step(acts[0, r], want_stats=True)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for k in range(rounds):
                for r, env in enumerate(envs):
                    env.step(acts[k, r], want_stats=True)
        g.replay()
        side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(replays):
            g.replay()
        e1.record(side)
        side.synchronize()
    torch.cuda.current_stream(device).wait_stream(side)
    steps = replays * rounds * replicas
    ms = e0.elapsed_time(e1) / steps
    touched = replicas * E * alg_bytes_per_env_step(kind, cfg, envs[0].W, envs[0].L)
    ws = sum(e._arena.numel() + e.obs.numel() * 4 for e in envs)
    out = {"value": E * N / (ms * 1e-3), "unit": UNIT, "us_per_step": 1e3 * ms, "steps": steps,
           "l2": "%d env batches round-robin: %.0f MB touched per round (algorithmic bytes; whole sectors are more) "
                 "> 126 MB L2, working set %.0f MB, no flush" % (replicas, touched / 1e6, ws / 1e6)}
    del envs
    torch.cuda.empty_cache()
    return out


def load_traffic(workload):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as f:
            return json.load(f).get(workload)
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=30)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c5", choices=sorted(WORKLOADS))
    ap.add_argument("--envs-per-gpu", type=int, default=0)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--sample-seconds", type=float, default=0.0,
                    help="--impl reference: seconds of stepping per sample (default: sized for a ~2 minute run)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
