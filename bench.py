#!/usr/bin/env python
"""bench.py — simulated env-µs/s of the fused WireEDMEnv step on N B200s (BASELINE.json metric).

A "step" is one control step: the fused kernel advances every environment through
servo_interval = 1000 µs of physics (1000 reference `WireEDMEnv.step()` calls per env).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs-per-gpu E] [--impl reference]

N > 1 is launched by torchrun (one rank per GPU, NCCL); invoked directly with --gpus > 1 it
re-executes itself under torch.distributed.run.  Rank 0 prints ONE JSON line.

Keys beyond the base contract:
  roofline        the BINDING bound of the step kernel: FP32 issue (algorithmic flop) against the
                  live-measured FFMA peak, plus `issue_frac` = warp-instructions per env-µs (committed SASS
                  histogram, profiles/ncu_traffic.json) x rate / (SMs x 4 schedulers x SM clock)
  roofline_hbm    the non-binding HBM side (state read once / written once per launch)
  cpu_baseline    the C port of the reference step on all host cores (kind "port")
  cpu_baseline_ref the reference's OWN WireEDMEnv.step (oracle/_ref, installed by oracle/fetch_ref.sh), one core and
                  all cores (kind "reference")
  cfg1            BASELINE configs[0]: one env, action_space.sample(), 1000 x 1-µs step(): drop-in vs reference
  cfg4            (N > 1) BASELINE configs[3]: 8192 envs per GPU with the stats all-reduce every 10 steps
  sustained       the same workload for >= 2 s with the clock sampler inside
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "simulated env-us/sec (batched WireEDMEnv.step)"
UNIT = "env-us/s"
WORKLOAD = ("cfg2: README default WireEDMEnv (brass 0.25 mm wire, 15 mm workpiece, 800 um target, "
            "N=375 wire segments), quickstart constant action (servo 0.1, 80 V, I5, ON 3 / OFF 80), "
            "on-device Philox draws")
README_CFG = dict(workpiece_height=15.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=800.0)
ACTION = (0.1, 80.0, 5, 3.0, 80.0)
N_SEG = 375
FLOP_PER_ENV_US = 13 * N_SEG + 60        # 13 flop/segment + ~60 scalar flop
DTYPE = "f64 scalars + f32 wire field"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--envs-per-gpu", type=int, default=4096)
    ap.add_argument("--impl", default="sparc_b200", choices=["sparc_b200", "reference"])
    ap.add_argument("--cpu-baseline-seconds", type=float, default=10.0)
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the CPU legs, cfg1 and the sustained run")
    ap.add_argument("--config", default="cfg2", choices=["cfg2", "cfg3", "cfg3_readme", "cfg5"],
                    help="cfg2 (default, the metric's config) | cfg3 position mode + P gap controller, 16384 envs | "
                         "cfg5 thick workpiece, fine segments, temperature-field logging")
    ap.add_argument("--segment-len", type=float, default=0.05, help="cfg5: wire segment length [mm]")
    ap.add_argument("--log-every", type=int, default=10, help="cfg5: sample the temperature field every N us")
    ap.add_argument("--host-controller", action="store_true",
                    help="cfg3: evaluate the gap controller with torch ops between launches instead of in the kernel latch")
    return ap.parse_args()


def config_dict(workload, E, world, n_seg, launch=None, extra=None):
    d = {"workload": workload, "envs_per_gpu": E, "n_envs": world * E, "us_per_step": 1000, "n_segments": n_seg,
         "l2": "flushed between timed steps (256 MiB fill)",
         "parallelism": f"env-sharded x{world}, NCCL all-reduce of fp64[8] stats every 10 steps",
         "launch": launch}
    if extra:
        d.update(extra)
    return d


# ------------------------------------------------------------------------------------------------
# CPU arms
# ------------------------------------------------------------------------------------------------
def cpu_oracle_throughput(target_seconds: float):
    """Times oracle/wedm_oracle.c (the pinned C restatement of the reference's CPU path) on every
    host core.  Bounded sample: 2 envs per core, horizon sized from a short probe."""
    from oracle import oracle as O

    cores = os.cpu_count() or 1
    p = O.make_params(**{k: v for k, v in README_CFG.items() if k != "wire_material"})
    t0 = time.perf_counter()
    done, _ = O.bench(p, ACTION, cores, 20000, cores, seed=1)
    probe = time.perf_counter() - t0
    rate = done / max(probe, 1e-6)
    n_envs = 2 * cores
    n_us = int(max(20000, min(5_000_000, rate * target_seconds / n_envs)))
    t0 = time.perf_counter()
    done, stats = O.bench(p, ACTION, n_envs, n_us, cores, seed=2)
    dt = time.perf_counter() - t0
    return dict(value=done / dt, unit=UNIT, cores=cores, kind="port",
                sample=f"{n_envs} envs x {n_us} us of the same workload on {cores} host threads "
                       f"(C restatement of the reference step, {dt:.1f} s)"), dt, done


def reference_throughput(seconds: float):
    """The reference's own WireEDMEnv.step (oracle/ref_bench.py): one core and all cores.  Returns the
    `cpu_baseline_ref` dict, or one that says why it could not be measured."""
    from oracle import ref_bench as RB

    ok, why = RB.available()
    if not ok:
        return {"kind": "reference", "unavailable": why}
    try:
        r = RB.time_reference(seconds)
        r["unit"] = UNIT
        r["value"] = r["all_cores"]
        return r
    except Exception as e:  # pragma: no cover - depends on the box
        return {"kind": "reference", "unavailable": f"{e.__class__.__name__}: {e}"}


def run_reference_arm(args):
    """The CPU arm the driver times beside the GPU arm: the reference's own env.step on all host cores
    (oracle/_ref) when it can be imported on this box, else the pinned C port; the other one rides along."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    steps = max(1, args.steps)
    budget = 100.0                                   # seconds for the whole run
    ref = reference_throughput(min(30.0, 0.45 * budget))
    per_step = min(6.0, 0.45 * budget / (steps + 1))
    cpu_oracle_throughput(0.5)                       # warm-up
    tot_t, tot_done, last = 0.0, 0, None
    for _ in range(steps):
        last, dt, done = cpu_oracle_throughput(per_step)
        tot_t += dt
        tot_done += done
    port_value = tot_done / tot_t
    last["value"] = port_value
    if "unavailable" in ref:
        value, base, note = port_value, last, "reference not importable here: value is the C port (" + ref["unavailable"] + ")"
    else:
        value, base, note = ref["all_cores"], dict(ref), "value is the reference's own WireEDMEnv.step on all host cores"
    E = args.envs_per_gpu
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": config_dict(WORKLOAD, E, max(1, args.gpus), N_SEG,
                              launch={"lanes_per_env": 32, "cells_per_lane": 12, "block_threads": 64, "blocks": (E + 1) // 2}),
        "cpu_baseline": base, "cpu_baseline_port": last, "cpu_baseline_ref": ref, "note": note,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out))
    return 0


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
                pw.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "power_w_max": max(pw) if pw else None, "reasons": sorted(reasons)}


def cfg1_dropin(dev):
    """BASELINE configs[0] on the drop-in surface: ONE env, reset(seed=0), 1000 x step(action_space.sample()),
    every call a 1-µs launch through the C-ABI with host buffers (what a user of the reference's API gets)."""
    from sparc_b200 import EnvironmentConfig, WireEDMEnv

    env = WireEDMEnv(config=EnvironmentConfig(**README_CFG), device=dev)
    env.reset(seed=0)
    env.action_space.seed(0)
    for _ in range(50):                              # warm-up: first launches, pinned buffers
        env.step(env.action_space.sample())
    env.reset(seed=0)
    actions = [env.action_space.sample() for _ in range(1000)]
    t0 = time.perf_counter()
    for a in actions:
        env.step(a)
    dt = time.perf_counter() - t0
    return {"us_per_call": 1e6 * dt / 1000, "calls": 1000, "time_after": int(env.state.time),
            "api": "sparc_b200.WireEDMEnv.step: one sparc_step_host(n_us=1) per call (H2D action, 1 kernel, D2H row, sync)"}


def run_gpu_arm(args):
    import torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        port = os.environ.get("MASTER_PORT", "29517")
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", port, os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)

    import torch.distributed as dist
    from sparc_b200 import build as sb_build
    if not os.environ.get("SPARC_LIB_PATH"):    # an explicitly named library (A/B experiments) is used as it is
        sb_build.build()
    from sparc_b200 import params as P
    from sparc_b200 import sharding
    from sparc_b200.engine import Engine, measure_fp32_peak

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # stdout carries exactly one JSON line: everything else written to fd 1 during the run (NCCL's
    # "NCCL version ..." banner comes from C code) is sent to stderr; the line itself goes to the saved fd
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    E = args.envs_per_gpu
    si = 1000
    workload, action, controller, log_every, kernel_ctrl = WORKLOAD, ACTION, None, 0, None
    sign = 1.0
    if args.config == "cfg2":
        prm = P.derive_params(P.EnvironmentConfig(**README_CFG), "position")
        assert prm.n_segments == N_SEG
    elif args.config in ("cfg3", "cfg3_readme"):
        # BASELINE cfg 3: position mode, proportional gap controller from wp - wire once per control step
        # (README.md:82-99; stabilising sign of run_simulation.py:28-33), evaluated INSIDE the kernel at the
        # action latch (sparc_set_controller) unless --host-controller asks for the torch-op version
        if "--envs-per-gpu" not in " ".join(sys.argv):
            E = 16384
        prm = P.derive_params(P.EnvironmentConfig(**README_CFG), "position")
        sign = 1.0 if args.config == "cfg3" else -1.0
        where = "torch ops between launches" if args.host_controller else "in the kernel's action latch"
        workload = (f"cfg3: position mode, P gap controller servo = clip({'+' if sign > 0 else '-'}0.1*(gap-25), -1, 1) "
                    f"evaluated {where} each control step, 80 V, I5, ON 3 / OFF 80, README default geometry, on-device Philox")
        if args.host_controller:
            def controller(engine):
                gap = engine["wp"] - engine["wire"]
                servo = torch.clamp(sign * 0.1 * (gap - 25.0), -1.0, 1.0).to(torch.float32)
                engine.set_actions(servo, 80.0, 5, 3.0, 80.0)
        else:
            kernel_ctrl = dict(kind=1, desired_gap=25.0, gain=sign * 0.1, limit=1.0)
    else:
        # BASELINE cfg 5: 50 mm workpiece, fine segments, I13 ON 3 / OFF 30, full field logging
        if "--envs-per-gpu" not in " ".join(sys.argv):
            E = 1024
        cfg5 = dict(workpiece_height=50.0, wire_diameter=0.25, wire_material="brass", target_cutting_distance=800.0)
        prm = P.derive_params(P.EnvironmentConfig(**cfg5), "position",
                              wire=P.WireModuleParameters(segment_len=args.segment_len))
        action = (0.0, 80.0, 13, 3.0, 30.0)
        log_every = args.log_every
        workload = (f"cfg5: 50 mm workpiece, segment_len {args.segment_len} mm (N={prm.n_segments}), I13 ON 3 / OFF 30, "
                    f"temperature field of every env logged every {log_every} us, on-device Philox")
    n_seg = int(prm.n_segments)
    flop_per_env_us = 13 * n_seg + 60
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def make_engine(envs):
        lo, _ = sharding.shard_range(world * envs, world, rank)      # contiguous global env ids per rank
        eng = Engine(envs, prm, device=dev, seed=20261104, env_id_base=lo)
        eng.set_actions(*action)
        if kernel_ctrl is not None:
            eng.set_controller(**kernel_ctrl)
        return eng

    def device_arm(eng, envs, K, warm, tlog=None, sampler=None):
        """K timed control steps with the state resident; returns (value, ms per step list, launches)."""
        def one_step():
            if controller is not None:
                controller(eng)
            if tlog is not None:
                eng.step_logged(si, log_every, out=tlog)
            else:
                eng.step(si)
        for _ in range(warm):
            one_step()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        launches = 0
        for k in range(K):
            flush.fill_(k & 0xFF)                       # evict the state from L2 (not timed)
            ev[k][0].record()
            one_step()
            launches += 2                               # mechanics pre-pass + fused step kernel
            if world > 1 and (k + 1) % 10 == 0 and not os.environ.get("SPARC_BENCH_NO_ALLREDUCE"):   # cfg 4: episode stats all-reduce every 10 control steps
                dist.all_reduce(eng.reduce_stats())
                launches += 1
            ev[k][1].record()
        barrier()
        step_ms = [a.elapsed_time(b) for a, b in ev]
        total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
        per_rank = [float(total_ms.item()) / K]
        if world > 1:
            allr = [torch.zeros_like(total_ms) for _ in range(world)]
            dist.all_gather(allr, total_ms)
            per_rank = [float(t.item()) / K for t in allr]
            dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        total_s = float(total_ms.item()) * 1e-3
        device_arm.per_rank_ms = per_rank
        return world * envs * si * K / total_s, step_ms, launches, total_s

    eng = make_engine(E)
    tlog = torch.zeros(E, si // log_every, n_seg, dtype=torch.float32, device=dev) if log_every else None

    # ---- device-resident arm ------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:                                   # one poller per box: eight of them perturb an 8-GPU run
        sampler.start()                             # samples cover warm-up + timed region (all under load)
    t_s = time.perf_counter()
    while rank == 0 and len(sampler.lines) < 3 and time.perf_counter() - t_s < 3.0:   # nvidia-smi needs ~0.2 s for its first line:
        eng.step(si)                                # keep the GPU under the same load until it reports,
        if eng["time"][0].item() > 150_000:         # so that even a 10-ms timed region has clock samples
            eng.reset(); eng.set_actions(*action)
    eng.reset()
    eng.set_actions(*action)
    K = args.steps
    t_wall0 = time.perf_counter()
    value, step_ms, launches, total_s = device_arm(eng, E, K, max(args.warmup, 3), tlog)
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    per_rank_ms = list(device_arm.per_rank_ms)

    # ---- end-to-end arm: C-ABI call with host buffers (H2D actions, D2H observation) ----------
    # Same workload as the device arm: cfg3 evaluates the gap controller on the HOST from the returned
    # observation (the way a reference user would); cfg5 also brings the logged field to pinned host
    # memory (through a <= 1 GiB pinned staging buffer, every byte crosses PCIe).
    from sparc_b200 import _native as N
    ha = eng.host_actions
    ha["servo"].fill_(action[0]); ha["target_voltage"].fill_(action[1]); ha["current_mode"].fill_(action[2])
    ha["on_time"].fill_(action[3]); ha["off_time"].fill_(action[4])
    d2h_extra = 0
    stage = None
    if tlog is not None:
        flat = tlog.view(-1)
        chunk = min(flat.numel(), (1 << 30) // 4)
        stage = torch.empty(chunk, dtype=torch.float32, pin_memory=True)
        d2h_extra = flat.numel() * 4

    servo_np = ha["servo"].numpy()                          # numpy view of the pinned action buffer
    if kernel_ctrl is not None:
        eng.set_controller(kind=0)                          # e2e: the controller runs on the host, from the observation

    def e2e_step(k, obs_prev):
        if args.config.startswith("cfg3") and obs_prev is not None:
            np.clip(sign * 0.1 * (obs_prev[0] - 25.0), -1.0, 1.0, out=servo_np)
        else:
            servo_np.fill(action[0] + 1e-4 * (k % 3))       # host-side action of this step
        if tlog is None:
            return eng.step_host(si).numpy()                # syncs; obs is pinned host memory (zero-copy view)
        eng.set_actions(ha["servo"], ha["target_voltage"], ha["current_mode"], ha["on_time"], ha["off_time"])
        eng.step_logged(si, log_every, out=tlog)
        for o in range(0, flat.numel(), chunk):
            n = min(chunk, flat.numel() - o)
            stage[:n].copy_(flat[o:o + n], non_blocking=True)
        return eng.observe().cpu().numpy()                  # syncs

    # same simulated window as the device arm: fresh episode, same number of warm-up control steps
    # (the workload drifts as the gap closes, so the two arms must not run back to back on one state)
    eng.reset()
    eng.set_actions(*action)
    obs = None
    for k in range(max(args.warmup, 3)):
        obs = e2e_step(k, obs)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    obs_sum = 0.0
    for k in range(K):
        obs = e2e_step(k, obs)
        obs_sum += float(obs[3].sum())
    e1.record()
    barrier()
    e2e_ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_value = world * E * si * K / (float(e2e_ms.item()) * 1e-3)

    total_stats = eng.reduce_stats().clone()
    if world > 1:
        dist.all_reduce(total_stats)
    ts = total_stats.cpu().tolist()
    launch_info = eng.launch_info()
    bytes_per_launch = eng.bytes_per_control_step()

    # ---- extra: BASELINE configs[3] at its stated size (N > 1 only) -----------------------------
    cfg4 = None
    if world > 1 and args.config == "cfg2":
        del eng
        torch.cuda.empty_cache()
        eng4 = make_engine(8192)
        v4, ms4, _, _ = device_arm(eng4, 8192, min(K, 50), 5)
        st4 = eng4.reduce_stats().clone()
        dist.all_reduce(st4)
        s4 = st4.cpu().tolist()
        cfg4 = {"value": v4, "unit": UNIT, "n_envs": world * 8192, "envs_per_gpu": 8192, "steps": min(K, 50),
                "ms_per_step": sum(ms4) / len(ms4),
                "stats_all_reduced": {"sparks": s4[0], "short_us": s4[1], "crater_um3": s4[3], "env_us": s4[5],
                                      "broken": s4[6], "target_reached": s4[7]},
                "workload": "cfg4: cfg2 settings, 8192 envs per GPU, NCCL all-reduce of the fp64[8] episode stats every 10 control steps"}
        eng = eng4

    # ---- extra: sustained run (>= 2 s, clocks sampled inside) -----------------------------------
    sustained = None
    if not args.no_cpu_baseline and args.config == "cfg2":
        eng.reset()
        eng.set_actions(*action)
        envs_s = eng.n_envs
        n_s = max(200, int(2.2 / max(1e-4, sum(step_ms) / len(step_ms) * 1e-3)))
        smp = ClockSampler(local_rank)
        if rank == 0:
            smp.start()
        for _ in range(5):
            eng.step(si)
        barrier()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for k in range(n_s):
            eng.step(si)
            if (k + 1) % 100 == 0:                          # keep the episode in the measured regime
                eng.reset()
        s1.record()
        barrier()
        sus_ms = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(sus_ms, op=dist.ReduceOp.MAX)
        sustained = {"value": world * envs_s * si * n_s / (float(sus_ms.item()) * 1e-3), "unit": UNIT, "steps": n_s,
                     "seconds": float(sus_ms.item()) * 1e-3, "envs_per_gpu": envs_s, "l2": "not flushed (back-to-back launches)",
                     "clocks": smp.stop() if rank == 0 else None}

    if rank == 0:
        fp32_peak = measure_fp32_peak(dev)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        hbm_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
        traffic, inst_per_env_us = None, None
        try:    # dram bytes and executed warp-instructions of one `ncu --set full` capture of this launch
            tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
            key = f"{args.config}:{E}" + (f":{n_seg}:{log_every}" if log_every else "")
            if key in tr:
                traffic = tr[key]["dram_bytes_read"] + tr[key]["dram_bytes_write"]
                inst_per_env_us = tr[key].get("warp_inst_per_env_us")
        except Exception:
            pass
        kern_ms = sum(step_ms) / K
        ach_gbs = bytes_per_launch / (kern_ms * 1e-3) / 1e9
        per_gpu_rate = E * si / (kern_ms * 1e-3)
        ach_tflops = per_gpu_rate * flop_per_env_us / 1e12
        if tlog is not None:
            bytes_per_launch += tlog.numel() * 4
            ach_gbs = bytes_per_launch / (kern_ms * 1e-3) / 1e9
        sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
        f_sm = (clocks.get("sm_mhz") or peaks.get("sm_max_mhz") or 1965.0) * 1e6
        issue_frac = (inst_per_env_us * per_gpu_rate / (sm_count * 4 * f_sm)) if inst_per_env_us else None
        fp32_roof = {"bound": "fp32_issue", "achieved": ach_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
                     "frac": ach_tflops / fp32_peak if fp32_peak else None, "traffic": traffic,
                     "flop_per_env_us": flop_per_env_us, "warp_inst_per_env_us": inst_per_env_us, "issue_frac": issue_frac,
                     "peak_source": "measured live: dense FFMA micro-kernel (sparc_measure_fp32_peak)",
                     "note": "nothing here is a contraction (no tensor cores); `achieved` counts the ALGORITHMIC flop of SURVEY 8d "
                             "(13 per segment + 60), `issue_frac` the instruction slots the kernel really uses "
                             "(warp-instructions per env-us from the committed ncu capture x rate / (SMs x 4 x SM clock))"}
        hbm_roof = {"bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
                    "traffic": traffic, "peak_source": hbm_src, "bytes_per_launch": bytes_per_launch,
                    "note": "state is read once and written once per 1000-us launch"}
        hbm_bound = tlog is not None                      # full-field logging is the HBM-bound configuration
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3),
            "ms_per_step": 1e3 * total_s / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": DTYPE, "data": "synthetic",
            "config": config_dict(workload, E, world, n_seg, launch_info),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 5 * 4 * E,
                    "d2h_bytes_per_step": N.N_OBS * 4 * E + d2h_extra,
                    "api": ("sparc_step_host (C-ABI, pinned host action/observation buffers)" if tlog is None else
                            "sparc_step_logged + D2H of the logged field to pinned host memory")},
            "gpu_launches": launches,
            "per_rank_ms_per_step": per_rank_ms,
            "clocks": clocks,
            "roofline": hbm_roof if hbm_bound else fp32_roof,
            ("roofline_fp32" if hbm_bound else "roofline_hbm"): fp32_roof if hbm_bound else hbm_roof,
            "stats": {"sparks": ts[0], "short_us": ts[1], "env_us": ts[5], "broken": ts[6], "target_reached": ts[7],
                      "sparks_per_env_ms": ts[0] / max(ts[5], 1.0) * 1000.0},
            "wall_s_timed_region": t_wall,
        }
        if cfg4 is not None:
            out["cfg4"] = cfg4
        if sustained is not None:
            out["sustained"] = sustained
        if not args.no_cpu_baseline and world == 1:
            cb, _, _ = cpu_oracle_throughput(args.cpu_baseline_seconds)
            out["cpu_baseline"] = cb
            ref = reference_throughput(args.cpu_baseline_seconds)
            out["cpu_baseline_ref"] = ref
            try:
                c1 = {"sparc_b200": cfg1_dropin(dev)}
                from oracle import ref_bench as RB
                c1["reference"] = RB.time_reference_cfg1() if RB.available()[0] else {"unavailable": RB.available()[1]}
                c1["workload"] = ("cfg1: README default env, reset(seed=0), 1000 x step(action_space.sample()) — one call = "
                                  "one us; us of wall time per call")
                out["cfg1"] = c1
            except Exception as e:  # pragma: no cover
                out["cfg1"] = {"error": f"{e.__class__.__name__}: {e}"}
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
