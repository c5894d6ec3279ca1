#!/usr/bin/env python
"""Benchmark of the per-env step path (BASELINE.json metric: env-steps/s, Go1 rough terrain, 187-point
height scan, full reward set) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs E] [--impl ours|reference]

One "step" = one LeggedRobot.step() without physics: action clip, position-target control for the
`decimation` substeps, and the fused post_physics_step kernel, over E envs per GPU of synthetic
simulator state (legged_gym_b200/utils/synthetic.py).  Prints ONE JSON line (rank 0).

  value     device-resident: the step is replayed from CUDA graphs over 8 pre-generated state frames
            that live in HBM (8 x 29 MB of inputs at 65536 envs, larger than L2).
  e2e       the same step through the public API with HOST buffers: every step copies that step's
            simulator state + actions from pinned host memory and reads obs / rewards / resets back.
  roofline  the kernels of post_physics_step (two launches; three with LGK_CHAIN=3) timed as one unit: algorithmic bytes per call / CUDA-event time
            (for --scenario go1_rough_actuator: the tensor-bound actuator MLP kernel).
  cpu_baseline  the reference's own, unmodified LeggedRobot.step() (staged into oracle/_ref by oracle/make_ref.sh,
            imported over the tensor-only simulator stand-in) timed on this box's host cores on a bounded sample of
            the same workload; the oracle port of it only where the staged copy is missing (kind says which).
  torch_cuda    the same reference code on CUDA torch, the same-box competitor.

`--impl reference` times that CPU implementation alone on the same config.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

SCENARIO = "go1_rough_full"
NUM_FRAMES = 8
METRIC = "env_steps_per_sec"
UNIT = "env-steps/s"


def workload_name(envs_per_gpu, decimation=4):
    """The one workload both arms name (BASELINE.json: Go1 rough, 4096-65536 envs/GPU)."""
    if SCENARIO != "go1_rough_full":      # --scenario: another BASELINE.json config, e.g. configs[2] (actuator net)
        return (f"scenario {SCENARIO} (legged_gym_b200/utils/scenarios.py), {envs_per_gpu} envs/GPU, "
                f"x{decimation} control substeps + fused post_physics_step")
    return (f"Go1 rough terrain, 187-point height scan, full reward set (20 terms + termination), terrain "
            f"curriculum, {envs_per_gpu} envs/GPU, position control x{decimation} substeps + fused post_physics_step")


def config_dict(envs_per_gpu, world, decimation=4, num_obs=235, reward_terms=21):
    """`config` of the JSON line: the same dict from both arms (the driver compares them key by key)."""
    return {"workload": workload_name(envs_per_gpu, decimation), "envs_per_gpu": envs_per_gpu,
            "envs_total": envs_per_gpu * world, "num_obs": num_obs, "reward_terms": reward_terms,
            "rng": "ours: in-kernel Philox4x32-7; reference: torch CPU generator",
            "l2": f"inputs cycle through {NUM_FRAMES} state frames resident in HBM "
                  f"({NUM_FRAMES * envs_per_gpu * 448 / 1e6:.0f} MB) -- larger than L2 at the default size",
            "parallelism": f"env-sharded x{world}, terrain replica per GPU, NCCL all-reduce of episode statistics "
                           f"inside the timed region"}


def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:   # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    def __init__(self, index=0):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:   # noqa: BLE001
            self.nv = None

    def _one(self):
        nv = self.nv
        self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
        except Exception:   # noqa: BLE001
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake", 0x2: "app_clocks_setting"}
        for bit, n in names.items():
            if r & bit:
                self.reasons.add(n)

    def _run(self):
        while not self._stop.is_set():
            try:
                self._one()
            except Exception:   # noqa: BLE001
                break
            time.sleep(0.002)

    def __enter__(self):
        if self.nv is not None and self._thr is None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def mark(self):
        """The timed region starts here: summary() reports the samples from this point on when there are at least
        three of them (a 20-step region is shorter than one sampling period), else the settle loop's as well."""
        self._mark = len(self.samples)

    def __exit__(self, *a):
        self._stop.set()
        if self._thr is not None:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        timed = self.samples[getattr(self, "_mark", 0):]
        s = sorted(timed if len(timed) >= 3 else self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "samples_in_timed_region": len(timed)}


def _dist_setup(n_gpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(0)
        local = 0
    return world, rank, local


def _build_env(envs, device, rank, world):
    import helpers

    def extra(cfg):
        cfg.env.shard_rank, cfg.env.shard_world = rank, world
        cfg.terrain.materialize_heights = False     # measured_heights stays in registers (algorithmic bytes)
        cfg.env.envs_per_cta = int(os.environ.get("LGK_ENVS_PER_CTA", "0"))
    env = helpers.make_env(SCENARIO, envs, device=device, seed=1 + rank, extra=extra)
    g = torch.Generator().manual_seed(99 + rank)
    env.episode_length_buf = torch.randint(0, int(env.max_episode_length), (envs,), generator=g).to(device)
    return env


def _bytes_per_env(env):
    from legged_gym_b200.utils.scenarios import algorithmic_bytes_per_env_step
    return algorithmic_bytes_per_env_step(env.num_dof, env.num_bodies, len(env.feet_indices), env._num_sum_rows,
                                          env.num_obs)


def run_ours(args):
    world, rank, local = _dist_setup(args.gpus)
    device = f"cuda:{local}"
    from legged_gym_b200 import _lgk
    from legged_gym_b200.utils import synthetic
    from legged_gym_b200.utils.distributed import ShardedStats
    lib = _lgk.load()
    env = _build_env(args.envs, device, rank, world)
    N = env.num_envs
    frames = synthetic.make_frames(env, NUM_FRAMES, seed=1234 + 1000 * 2 + rank, device=device)
    # the statistics all-reduce (the path's only collective) must fire inside any timed region, however short
    # production cadence is every 24 steps; a shorter timed region gets exactly one all-reduce, half-way through it
    # (see `phase` below).  Measured at 2 GPUs, 20 steps: each all-reduce in the window costs ~45 us -- its own CTA
    # and the three small kernels around it take SM slots from the chain's grids for one step
    stats_interval = min(24, max(1, args.steps))
    stats = ShardedStats(env, interval=stats_interval) if world > 1 else None
    n_collectives = [0]
    stream = torch.cuda.Stream(device)
    torch.cuda.synchronize()

    # ---- device-resident path: one CUDA graph per frame ------------------------------------------
    graphs, launches_per_step = [], 0
    with torch.cuda.stream(stream):
        env.bind_sim_tensors(**{k: frames[0][k] for k in ("root_states", "dof_state", "contact_forces", "torques")})
        for _ in range(3):      # eager warm-up (loads kernels, seeds the device step counter)
            env.step(frames[0]["actions"])
        stream.synchronize()
        for k in range(NUM_FRAMES):
            fr = frames[k]
            env.bind_sim_tensors(**{n: fr[n] for n in ("root_states", "dof_state", "contact_forces", "torques")})
            g = torch.cuda.CUDAGraph()
            l0 = lib.lgk_launch_count()
            with torch.cuda.graph(g, stream=stream):
                env.step(fr["actions"])
            launches_per_step = lib.lgk_launch_count() - l0
            graphs.append(g)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    # phase of the periodic all-reduce against the timed window: half an interval in, so that no all-reduce is the
    # window's very last event (its whole latency would then be exposed, which no steady-state step sees)
    phase = stats_interval // 2

    def replay(i):
        graphs[i % NUM_FRAMES].replay()
        if stats is not None and stats.after_step(i + phase) is not None:
            n_collectives[0] += 1

    with torch.cuda.stream(stream):
        # (with several ranks the warm-up runs at least two statistics all-reduces: the first one sets NCCL's
        # connections up, which is not a timed event)
        n_warm = max(args.warmup, 3) if stats is None else max(args.warmup, 2 * stats_interval)
        for i in range(n_warm):
            replay(i)
        # the clock sampler (NVML set-up, thread start) comes up BEFORE the ranks line up: started behind the barrier it
        # skewed the ranks' windows against each other by its own start-up time, and with an all-reduce inside the window
        # the early rank then waits for the late one.  It samples from the settle loop on (the same graph replays as
        # the timed region), marked below.
        sampler = ClockSampler(local)
        sampler.__enter__()
        t_end = time.time() + 0.3          # settle clocks under load before timing
        i = 0
        while time.time() < t_end:         # time-bounded: NO collective in here (the ranks run different counts)
            graphs[i % NUM_FRAMES].replay()
            i += 1
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with sampler as clocks:
            clocks.mark()
            n_collectives[0] = 0
            e0.record(stream)
            for i in range(args.steps):
                replay(i)
            e1.record(stream)
            if stats is not None:
                stats.wait()            # the side-stream all-reduces belong to the timed region
                e1 = torch.cuda.Event(enable_timing=True)
                e1.record(stream)
            collectives_timed = n_collectives[0]
            barrier()
        ms_total = e0.elapsed_time(e1)
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms_total], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    value = N * world * args.steps / (ms_total * 1e-3)
    env._dev_step_counter = -1      # graph replays advanced the device step counter: reseed it on the next eager step

    # ---- dominant kernel alone: events around the fused post-physics launch ---------------------
    import ctypes as C
    with torch.cuda.stream(stream):
        st = C.c_void_p(stream.cuda_stream)
        reps = max(20, min(args.steps, 200))
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
        kt = []
        for i in range(reps + 3):
            fr = frames[i % NUM_FRAMES]
            env.bind_sim_tensors(**{n: fr[n] for n in ("root_states", "dof_state", "contact_forces", "torques")})
            flush.zero_()                       # evict L2 between timed launches
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(stream)
            _lgk.check("step", lib.lgk_post_physics_step(C.byref(env._consts), C.byref(env._bufs), st))
            a1.record(stream)
            stream.synchronize()
            if i >= 3:
                kt.append(a0.elapsed_time(a1))
        del flush
    k_ms = sum(kt) / len(kt)
    bpe = _bytes_per_env(env)
    peak, peak_src = _peaks()
    achieved = bpe * N / (k_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "step_kernel_traffic.json")
    if os.path.isfile(tpath):
        try:
            traffic = json.load(open(tpath)).get(str(N))
        except Exception:   # noqa: BLE001
            traffic = None
    chain3 = os.environ.get("LGK_CHAIN", "")[:1] == "3"
    roofline = {"kernel": ("post_physics_step = lgk_scan_kernel + lgk_step_kernel + lgk_finish_kernel (LGK_CHAIN=3)" if chain3
                           else "post_physics_step = lgk_step_kernel + lgk_tail_kernel (height scan, reward "
                                "sums, reset tail, statistics)") +
                          " (chained with programmatic dependent launch; timed as one unit)",
                "bound": "hbm", "achieved": round(achieved, 1), "peak": peak,
                "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                "algorithmic_bytes_per_env_step": bpe, "kernel_ms": round(k_ms, 5), "peak_source": peak_src,
                "timing": "CUDA events around single lgk_post_physics_step calls, L2 flushed (256 MB memset) before each"}
    if getattr(env.cfg.control, "use_actuator_network", False) and hasattr(env, "actuator_advance"):
        # the dominant kernel of this scenario is the actuator MLP (decimation launches per step), tensor-bound
        roofline = {"post_physics_step": roofline, **_actuator_roofline(env, frames, stream)}

    # ---- end to end through the public API with host buffers --------------------------------------
    e2e = None
    if not args.no_e2e:
        chunks = args.e2e_chunks if (N % (args.e2e_chunks * 128) == 0 and N // args.e2e_chunks >= 8192) else 1
        if chunks > 1:
            try:
                e2e = run_e2e_chunked(frames, args, world, rank, device, chunks)
            except Exception as exc:   # noqa: BLE001  (fall back to the one-env path; say why)
                print(f"[bench] chunked e2e leg failed ({type(exc).__name__}: {exc}); falling back", file=sys.stderr)
                e2e = None
        if e2e is None:
            e2e = run_e2e(env, frames, args, world, device, stream)

    # ---- CPU baseline (rank 0, single-GPU runs only) ------------------------------------------------
    cpu = None
    torch_cuda = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference(args.envs, budget_s=args.cpu_budget)
        try:
            torch_cuda = torch_cuda_competitor(args.envs, device)
        except Exception as exc:   # noqa: BLE001
            torch_cuda = {"unavailable": f"{type(exc).__name__}: {exc}"[:200]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(ms_total / args.steps, 6), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(N, world, env.cfg.control.decimation, env.num_obs, env._num_sum_rows),
            "collectives_in_timed_region": collectives_timed, "stats_allreduce_interval": stats_interval if world > 1 else None,
            "e2e": e2e, "gpu_launches": int(launches_per_step * args.steps), "launches_per_step": int(launches_per_step),
            "roofline": roofline, "cpu_baseline": cpu, "torch_cuda": torch_cuda, "clocks": clocks.summary(),
            # workspace header word 5 (LgkWorkspace.scan_timeouts): warps that gave up waiting for a hand-over flag
            # (three-launch chain only; must be 0)
            "scan_timeouts": int(env._workspace.view(torch.int32)[5].item()),
        }
        print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


def run_e2e_chunked(frames, args, world, rank, device, chunks):
    """The host-buffer path with the GPU's envs held by `chunks` LeggedRobot instances (env-sharded exactly like ranks:
    global env ids, own Philox offsets), each stepping on a stream of its own: PCIe is the limiter of this path (the
    940 B observation row of every env crosses it every step), so the read-back of chunk k's results runs while chunk
    k + 1 computes and only the first chunk's kernels and the actions' upload stay exposed.  Every step still ends
    with a host sync on all chunks (the caller consumes the step's result before the next one)."""
    names = ("root_states", "dof_state", "contact_forces", "torques", "actions")
    sim_names = names[:4]
    N = frames[0]["actions"].shape[0]
    n = N // chunks
    envs = [_build_env(n, device, rank * chunks + c, world * chunks) for c in range(chunks)]
    host = [{k: fr[k].cpu().pin_memory() for k in names} for fr in frames]
    rows = lambda t, c: t.view(N, -1)[c * n:(c + 1) * n]          # chunk c's rows of a per-env tensor
    dev = [[{k: torch.empty_like(rows(frames[0][k], c)).contiguous() for k in names} for _ in range(2)] for c in range(chunks)]
    streams = [torch.cuda.Stream(device) for _ in range(chunks)]
    copy_stream = torch.cuda.Stream(device)
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    O = envs[0].obs_buf.shape[1]
    out_h = {"obs": torch.empty(N, O, dtype=torch.float).pin_memory(), "rew": torch.empty(N, dtype=torch.float).pin_memory(),
             "reset": torch.empty(N, dtype=torch.bool).pin_memory()}
    h2d = sum(host[0][k].numel() * host[0][k].element_size() for k in names)
    d2h = sum(v.numel() * v.element_size() for v in out_h.values())
    shape = {k: dev[0][0][k].shape for k in names}
    # one captured step() per chunk and input set (public API, graph-capturable)
    graphs = [[None, None] for _ in range(chunks)]
    for c, env in enumerate(envs):
        with torch.cuda.stream(streams[c]):
            for k in range(2):
                for nm in names:
                    dev[c][k][nm].copy_(rows(frames[0][nm], c).view(shape[nm]))
                env.bind_sim_tensors(**{nm: dev[c][k][nm] for nm in sim_names})
                env.step(dev[c][k]["actions"])
                streams[c].synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=streams[c]):
                    env.step(dev[c][k]["actions"])
                graphs[c][k] = g
    torch.cuda.synchronize()

    def prefetch(i):
        h = host[i % len(host)]
        with torch.cuda.stream(copy_stream):
            for c in range(chunks):
                for nm in sim_names:
                    dev[c][i % 2][nm].copy_(rows(h[nm], c).view(shape[nm]), non_blocking=True)
            ready[i % 2].record(copy_stream)

    def one(i):
        h = host[i % len(host)]
        for c, env in enumerate(envs):
            st = streams[c]
            with torch.cuda.stream(st):
                dev[c][i % 2]["actions"].copy_(rows(h["actions"], c).view(shape["actions"]), non_blocking=True)
                st.wait_event(ready[i % 2])
                graphs[c][i % 2].replay()
                out_h["obs"][c * n:(c + 1) * n].copy_(env.obs_buf, non_blocking=True)
                out_h["rew"][c * n:(c + 1) * n].copy_(env.rew_buf, non_blocking=True)
                out_h["reset"][c * n:(c + 1) * n].copy_(env.reset_buf, non_blocking=True)
        prefetch(i + 1)                 # the other input set: step i-1 finished with it before the last sync
        for st in streams:
            st.synchronize()            # the caller consumes the step's result before the next one

    prefetch(0)
    steps = max(10, min(args.steps, 200))
    for i in range(3):
        one(i)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    main = torch.cuda.current_stream(device)
    e0.record(main)
    for st in streams:
        st.wait_event(e0)
    for i in range(steps):
        one(i)
    for st in streams:
        main.wait_stream(st)
    e1.record(main)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    assert bool(torch.isfinite(out_h["obs"]).all()) and float(out_h["obs"].abs().sum()) > 0.0
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return {"value": round(N * world * steps / (ms * 1e-3), 1), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "steps": steps, "ms_per_step": round(ms / steps, 5), "chunks": chunks,
            "path": f"pinned host buffers -> LeggedRobot.step() -> pinned host buffers, host-synchronised every step; the "
                    f"GPU's envs are held by {chunks} env instances of {n} envs stepping on streams of their own (each "
                    f"step() captured once through the public API and replayed), so that chunk k's read-back runs under "
                    f"chunk k+1's kernels; next step's simulator state is copied in on a further stream meanwhile"}


def _actuator_roofline(env, frames, stream):
    """Go1 actuator substep (history kernel + tcgen05 MLP kernel) timed alone, against the tensor pipe: every product
    runs as three TF32 passes (layer 0's K = 30 padded to 32); TF32 peak = half the measured dense bf16 rate, the burst
    figure because the kernel is timed in isolation."""
    N = env.num_envs
    act = frames[0]["actions"]
    with torch.cuda.stream(stream):
        for _ in range(5):
            env.actuator_advance(act)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 50
        e0.record(stream)
        for _ in range(reps):
            env.actuator_advance(act)
        e1.record(stream)
        stream.synchronize()
    ms = e0.elapsed_time(e1) / reps
    pipe = 3 * 2 * (32 * 128 + 2 * 128 * 128) * 4 * N
    useful = 2 * (30 * 128 + 2 * 128 * 128 + 128 * 3) * 4 * N
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peak, src = float(pk["bf16_tflops"]) / 2, "measured (MEASURED_PEAKS.json bf16_tflops / 2: TF32 runs at half the bf16 rate; burst figure)"
    except Exception:   # noqa: BLE001
        peak, src = 2250.0 / 2, "fallback (B200_PROFILING.md dense bf16 2.25 PFLOP/s / 2)"
    ach = pipe / (ms * 1e-3) / 1e12
    return {"kernel": "lgk_act_mlp_umma_kernel (+ lgk_act_hist_kernel ahead of it, timed together)", "bound": "tensor",
            "achieved": round(ach, 1), "peak": round(peak, 1), "unit": "TFLOP/s", "frac": round(ach / peak, 4), "traffic": None,
            "kernel_ms": round(ms, 5), "peak_source": src,
            "flops_per_launch": {"tensor_pipe_3xTF32": pipe, "useful_fp32_equivalent": useful,
                                 "useful_tflops": round(useful / (ms * 1e-3) / 1e12, 1)},
            "timing": "CUDA events over 50 back-to-back substeps (same inputs; weights and inputs L2-resident)"}


def run_e2e(env, frames, args, world, device, stream):
    """Public-API path with host buffers: pinned host -> device copies of that step's simulator state and
    actions, env.step(), device -> pinned host read-back of observations, rewards and resets."""
    N = env.num_envs
    names = ("root_states", "dof_state", "contact_forces", "torques", "actions")
    host = [{k: fr[k].cpu().pin_memory() for k in names} for fr in frames]
    sim_names = names[:4]
    # two device input sets: the simulator state of step i+1 travels host -> device on a copy stream while
    # step i's results travel device -> host (PCIe is full duplex).  The actions of step i+1 depend on the
    # observations of step i in a real loop, so they are copied only after the read-back has completed.
    dev = [{k: torch.empty_like(frames[0][k]) for k in names} for _ in range(2)]
    copy_stream = torch.cuda.Stream(device)
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    out_h = {"obs": torch.empty(env.obs_buf.shape, dtype=torch.float).pin_memory(),
             "rew": torch.empty(N, dtype=torch.float).pin_memory(),
             "reset": torch.empty(N, dtype=torch.bool).pin_memory()}
    h2d = sum(host[0][k].numel() * host[0][k].element_size() for k in names)
    d2h = sum(v.numel() * v.element_size() for v in out_h.values())

    def prefetch(i):
        h, d = host[i % len(host)], dev[i % 2]
        with torch.cuda.stream(copy_stream):
            for k in sim_names:
                d[k].copy_(h[k], non_blocking=True)
            ready[i % 2].record(copy_stream)

    # env.step() on each of the two input sets, captured once through the public API (LeggedRobot.step() is
    # graph-capturable and sync-free, tests/test_gpu_parity.py): a replay costs one launch instead of the ~10 ctypes
    # calls of an eager step, whose host time sits on the critical path when every step ends in a host sync
    step_graphs = []
    zero_copy = bool(getattr(args, "e2e_zero_copy", False))
    if zero_copy:
        env.bind_output_tensors(obs_buf=out_h["obs"], rew_buf=out_h["rew"])
    if not args.e2e_eager:
        with torch.cuda.stream(stream):
            for k in range(2):
                env.bind_sim_tensors(**{n: dev[k][n] for n in sim_names})
                env.step(dev[k]["actions"])            # eager once: seeds the device step counter
                stream.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    env.step(dev[k]["actions"])
                step_graphs.append(g)
        torch.cuda.synchronize()

    def one(i):
        h, d = host[i % len(host)], dev[i % 2]
        d["actions"].copy_(h["actions"], non_blocking=True)
        stream.wait_event(ready[i % 2])
        if step_graphs:
            step_graphs[i % 2].replay()
            obs, rew, reset = env.obs_buf, env.rew_buf, env.reset_buf
        else:
            env.bind_sim_tensors(**{k: d[k] for k in sim_names})
            obs, _, rew, reset, _ = env.step(d["actions"])
        prefetch(i + 1)                 # the other input set: step i-1 finished with it before the last sync
        if not zero_copy:
            out_h["obs"].copy_(obs, non_blocking=True)
            out_h["rew"].copy_(rew, non_blocking=True)
        out_h["reset"].copy_(reset, non_blocking=True)
        stream.synchronize()            # the caller consumes the step's result before the next one

    prefetch(0)
    steps = max(10, min(args.steps, 200))
    with torch.cuda.stream(stream):
        for i in range(3):
            one(i)
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(steps):
            one(i)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    # where the time goes: each leg alone (events, 10 repetitions), this rank
    def leg(fn):
        with torch.cuda.stream(stream):
            fn(); stream.synchronize()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(stream)
            for _ in range(10):
                fn()
            a1.record(stream)
            stream.synchronize()
        return a0.elapsed_time(a1) / 10

    def leg_h2d():
        for k in names:
            dev[0][k].copy_(host[0][k], non_blocking=True)

    def leg_d2h():
        out_h["obs"].copy_(env.obs_buf, non_blocking=True)
        out_h["rew"].copy_(env.rew_buf, non_blocking=True)
        out_h["reset"].copy_(env.reset_buf, non_blocking=True)

    def leg_step():
        if step_graphs:
            step_graphs[0].replay()
        else:
            env.step(dev[0]["actions"])
    t_h2d, t_d2h, t_step = leg(leg_h2d), leg(leg_d2h), leg(leg_step)
    env._dev_step_counter = -1      # graph replays advanced the device step counter behind the host's copy
    breakdown = {"h2d_ms": round(t_h2d, 4), "h2d_gbs": round(h2d / t_h2d / 1e6, 1), "step_kernels_ms": round(t_step, 4),
                 "d2h_ms": round(t_d2h, 4), "d2h_gbs": round(d2h / t_d2h / 1e6, 1),
                 "sync_and_launch_ms": round(max(ms / steps - t_step - t_d2h, 0.0), 4),
                 "note": "per step the chain is actions H2D -> kernels -> D2H -> host sync; the simulator-state H2D of the "
                         "next step runs on a second stream under this step's D2H.  The limiter is the D2H of the "
                         "observation rows over PCIe"}
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return {"value": round(N * world * steps / (ms * 1e-3), 1), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "steps": steps, "ms_per_step": round(ms / steps, 5), "breakdown": breakdown,
            "path": "pinned host buffers -> LeggedRobot.step() ("
                    + ("captured once per input set through the public API and replayed as a CUDA graph" if step_graphs
                       else "eager")
                    + ") -> pinned host buffers, stream-synchronised every step; next step's simulator state is copied in "
                      "(second stream) while this step's results are copied out"}


def torch_cuda_competitor(envs, device):
    """The "same box" competitor: the reference's own op-by-op torch
    implementation of the path run on the GPU -- the unmodified reference over the simulator stand-in when it is here
    (oracle/_ref), else the oracle port of it -- timed with CUDA events, median of a few steps."""
    from legged_gym_b200.utils import scenarios, synthetic
    from oracle import reference_loader
    n, steps, warm = envs, 12, 4
    out = {"envs": n, "torch": torch.__version__}
    try:
        if not reference_loader.reference_available():
            raise RuntimeError("reference not staged (oracle/_ref)")
        task, mutate = scenarios.SCENARIOS[SCENARIO]
        env = reference_loader.make_reference_env(task, mutate, num_envs=n, device=device, seed=1)
        g = torch.Generator().manual_seed(99)
        env.episode_length_buf[:] = torch.randint(0, int(env.max_episode_length), (n,), generator=g).to(device)
        frames = [{k: v.to(device) for k, v in fr.items()} for fr in synthetic.make_frames(env, 2, seed=3234)]

        def one(i):
            synthetic.load_frame(env, frames[i % 2])
            a = frames[i % 2]["actions"].clone()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            env.step(a)
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1)
        out["kind"] = f"the reference's own LeggedRobot.step() on CUDA torch ({reference_loader.reference_kind()})"
    except Exception as exc:   # noqa: BLE001  (fall back to the port; say why)
        import helpers
        from oracle import legged_oracle as lo
        out["reference_on_cuda_failed"] = f"{type(exc).__name__}: {exc}"[:200]
        envc = helpers.make_env(SCENARIO, n, device="cpu", seed=1)
        g = torch.Generator().manual_seed(99)
        envc.episode_length_buf = torch.randint(0, int(envc.max_episode_length), (n,), generator=g)
        p = lo.OracleParams.from_env(envc, device=device)
        s = lo.OracleState.from_env(envc, p, device=device)
        frames = [{k: v.to(device) for k, v in fr.items()} for fr in synthetic.make_frames(envc, 2, seed=3234)]
        taps = {k: v.to(device) for k, v in synthetic.make_taps(envc, seed=5).items()}

        def one(i):
            fr = frames[i % 2]
            s.root_states.copy_(fr["root_states"]); s.dof_state.copy_(fr["dof_state"])
            s.contact_forces.copy_(fr["contact_forces"].view_as(s.contact_forces)); s.torques.copy_(fr["torques"])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            taps["noise"] = torch.rand(n, envc.num_obs, device=device)
            lo.step(p, s, fr["actions"], taps)
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1)
        out["kind"] = "oracle port of the reference on CUDA torch"
    ts = sorted([one(i) for i in range(steps + warm)][warm:])
    ms = ts[len(ts) // 2]
    out.update(ms_per_step=round(ms, 4), value=round(n / (ms * 1e-3), 1), unit=UNIT,
               note="median of 12 steps, CUDA events; includes the reference's own host syncs (nonzero() on the reset / "
                    "resample id lists)")
    return out


def cpu_reference(envs, budget_s=20.0, steps=None, warmup=1):
    """The reference's CPU implementation of the path (oracle port) on all host cores, bounded sample."""
    import helpers
    from legged_gym_b200.utils import synthetic
    from oracle import legged_oracle as lo
    from oracle import reference_loader
    torch.set_num_threads(os.cpu_count() or 1)
    # the reference's own code when it is here (oracle/_ref, staged by oracle/make_ref.sh; /root/reference in the
    # build container), else the oracle port of it
    use_ref = reference_loader.reference_available()
    timer = (lambda n, k, w: _time_reference(n, k, w, synthetic, reference_loader)) if use_ref else \
            (lambda n, k, w: _time_oracle(n, k, w, helpers, synthetic, lo))
    # size the sample from a short probe so the whole leg stays within the budget
    probe_n = min(envs, 4096)
    t_probe = timer(probe_n, 2, 1)
    per_env = t_probe / probe_n
    n = envs
    if steps is None:
        steps = 5
    while n > 1024 and per_env * n * (steps + warmup) > budget_s:
        n //= 2
    t = timer(n, steps, warmup)
    what = (f"the reference's own LeggedRobot.step() ({reference_loader.reference_kind()}, unmodified, over the "
            f"tensor-only simulator stand-in)") if use_ref else "oracle/legged_oracle.py"
    return {"value": round(n / t, 1), "unit": UNIT, "cores": torch.get_num_threads(), "kind": "reference" if use_ref else "port",
            "sample": f"{steps} steps x {n} envs of the same workload after {warmup} warm-up ({what}, "
                      f"torch {torch.__version__} CPU kernels)", "ms_per_step": round(t * 1e3, 3)}


_CPU_META = {"decimation": 4, "num_obs": 235, "reward_terms": 21}


def _time_reference(n, steps, warmup, synthetic, reference_loader):
    """The reference's unmodified task class on CPU torch: env.step(actions) through its public API, stock random
    draws, with a synthetic simulator frame written in place of gym.simulate()'s result before each step."""
    from legged_gym_b200.utils import scenarios
    task, mutate = scenarios.SCENARIOS[SCENARIO]
    env = reference_loader.make_reference_env(task, mutate, num_envs=n, seed=1)
    g = torch.Generator().manual_seed(99)
    env.episode_length_buf[:] = torch.randint(0, int(env.max_episode_length), (n,), generator=g)
    _CPU_META.update(decimation=int(env.cfg.control.decimation), num_obs=int(env.num_obs), reward_terms=len(env.episode_sums))
    frames = synthetic.make_frames(env, 2, seed=3234)
    ts = []
    for i in range(steps + warmup):
        fr = frames[i % 2]
        synthetic.load_frame(env, fr)
        a = fr["actions"].clone()
        t0 = time.perf_counter()
        env.step(a)
        ts.append(time.perf_counter() - t0)
    ts = sorted(ts[warmup:])
    return ts[len(ts) // 2]


def _time_oracle(n, steps, warmup, helpers, synthetic, lo):
    env = helpers.make_env(SCENARIO, n, device="cpu", seed=1)
    g = torch.Generator().manual_seed(99)
    env.episode_length_buf = torch.randint(0, int(env.max_episode_length), (n,), generator=g)
    p = lo.OracleParams.from_env(env)
    s = lo.OracleState.from_env(env, p)
    _CPU_META.update(decimation=int(env.cfg.control.decimation), num_obs=int(env.num_obs), reward_terms=int(env._num_sum_rows))
    frames = synthetic.make_frames(env, 2, seed=3234)
    taps = synthetic.make_taps(env, seed=5)
    ts = []
    for i in range(steps + warmup):
        fr = frames[i % 2]
        s.root_states.copy_(fr["root_states"]); s.dof_state.copy_(fr["dof_state"])
        s.contact_forces.copy_(fr["contact_forces"].view_as(s.contact_forces)); s.torques.copy_(fr["torques"])
        t0 = time.perf_counter()
        taps["noise"] = torch.rand(n, env.num_obs)      # the reference draws its noise every step (:231)
        lo.step(p, s, fr["actions"], taps)
        ts.append(time.perf_counter() - t0)
    ts = sorted(ts[warmup:])
    return ts[len(ts) // 2]


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.perf_counter()
    cpu = cpu_reference(args.envs, budget_s=max(30.0, min(150.0, 1.5 * (args.steps + args.warmup))),
                        steps=max(1, min(args.steps, 20)), warmup=max(1, min(args.warmup, 3)))
    n_sample = int(cpu["sample"].split(" x ")[1].split(" envs")[0])
    line = {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cpu["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args.envs, max(1, args.gpus), _CPU_META["decimation"], _CPU_META["num_obs"], _CPU_META["reward_terms"]),
            "cpu_sample_envs_per_step": n_sample,
            "cpu_baseline": cpu, "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0,
                                         "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": round(time.perf_counter() - t0, 1)}
    print(json.dumps(line))


def main():
    global SCENARIO
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--envs", type=int, default=65536, help="envs per GPU (BASELINE.json: 4096-65536)")
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="development: skip the host-buffer leg (e2e: null)")
    ap.add_argument("--e2e-eager", action="store_true", help="e2e leg with an eager step() instead of a captured one")
    ap.add_argument("--e2e-zero-copy", action="store_true",
                    help="e2e leg (one env): the kernels write observations and rewards straight into the pinned host buffers "
                         "(bind_output_tensors on mapped host memory) instead of a device->host copy after the step")
    ap.add_argument("--e2e-chunks", type=int, default=1,
                    help="e2e leg: the envs of a GPU as this many env instances stepping on streams of their own, so that "
                         "the read-back of one chunk's observations runs under the next chunk's kernels (1 = one env)")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    ap.add_argument("--scenario", default=SCENARIO,
                    help="parity-test scenario to time instead of the headline one, e.g. go1_rough_actuator "
                         "(BASELINE.json configs[2], with --envs 16384)")
    args = ap.parse_args()
    SCENARIO = args.scenario
    if os.environ.get("LGK_BENCH_WATCHDOG"):     # development: dump every thread's Python stack if the run stalls
        import faulthandler
        faulthandler.dump_traceback_later(float(os.environ["LGK_BENCH_WATCHDOG"]), exit=True)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
