#!/usr/bin/env python
"""bench.py -- throughput of the collision-probability hot path on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scene pr2test2|pr2custom]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[2], "MC convergence run"): Monte-Carlo rollouts of the committed 56-waypoint
plan (tests/golden/plan_committed.json = trajectory.dat / odometry.dat as MCSimulation.py sends them) with the
reference's noise parameters, Philox mode (every rollout owns its real-robot/EKF thread and its particle,
SURVEY.md A0), NO early exit (the reference moves and tests every particle at every waypoint,
MCSimulator.h:333-347), so executed rollout-steps == nominal rollout-steps.  One step = one runSimulation of
`--rollouts` rollouts PER GPU (weak scaling: rank r owns rollout ids [r*n, (r+1)*n), a disjoint Philox
subsequence per id) followed by the NCCL all-reduce of the hit counters on the same stream.

metric  = MC rollout-steps/s (a rollout-step = one pose update + one robot-vs-scene collision query).
value   = device-resident: plan uploaded once, timed region = K x (rollout kernel + all-reduce), CUDA events on
          the library's own stream, max over ranks.
e2e     = the same run through the plugin command interface (`SendCommand("runSimulation")`, host buffers in,
          P(collision) text out), host<->device copies inside the timed region.
roofline: the rollout kernel is FP32-pipe bound (state never leaves registers, zero HBM traffic per
          rollout-step), so `bound` is "fp32": achieved = algorithmic flops (647 per Philox
          rollout-step + 40 per SAT pair test, counted by the kernel) / kernel time; peak = the FFMA-chain
          microbenchmark of csrc/peaks_sm100.cuh measured in this same process (MEASURED_PEAKS.json holds HBM /
          bf16 peaks only; its values are quoted under roofline.file_peaks).
cpu_baseline / --impl reference: the reference's own estimator headers (oracle/_ref/libmcsim_ref.so, built
          from /root/reference where it lies) with the CPU SAT checker, timed on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
F_PHILOX_STEP, F_SAT, F_GMM_SAMPLE = 647, 40, 35  # algorithmic flops (DESIGN.md "Algorithmic work")
# algorithmic MUFU operations of one Philox rollout-step: 2 moves x 2 + 11 normals x 2.5 + controller 6 + 8 updates x 2
SFU_PHILOX_STEP = 2 * 2 + 11 * 2.5 + 6 + 8 * 2
METRIC, UNIT = "mc_rollout_steps_per_s", "rollout-steps/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scene", default="pr2test2", choices=["pr2test2", "pr2custom"])
    ap.add_argument("--rollouts", type=int, default=10 ** 8, help="rollouts per GPU per step")
    ap.add_argument("--e2e-rollouts", type=int, default=10 ** 7, help="rollouts per GPU per e2e step")
    ap.add_argument("--ref-particles", type=int, default=10000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exhaustive", action="store_true", help="time the rollout kernel with its broad phase off (every live pair at every step)")
    ap.add_argument("--no-extras", action="store_true", help="skip the GMM / early-exit / pr2custom side measurements")
    return ap.parse_args()


def load_inputs(scene):
    import pcoll_b200
    plan = pcoll_b200.load_plan_json(os.path.join(GOLD, "plan_committed.json"))
    sc = np.load(os.path.join(GOLD, "scene_%s.npz" % scene))
    return plan, dict(boxes=sc["boxes"], hulls=sc["hulls"])


# ---------------------------------------------------------------------------------------------------
# clocks: sampled with NVML while the timed region runs
class ClockSampler(threading.Thread):
    REASONS = (("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40),
               ("hw_power_brake_slowdown", 0x80), ("sw_power_cap", 0x4), ("sync_boost", 0x10))

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        while not self.stop_flag:
            try:
                if self.nv:
                    self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                    mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    for name, bit in self.REASONS:
                        if mask & bit:
                            self.reasons.add(name)
                else:
                    o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.active",
                                        "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout.split(",")
                    self.samples.append(int(o[0]))
                    self.max_mhz = int(o[1])
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        self.stop_flag = True
        self.join(timeout=2)
        s = sorted(self.samples)
        return dict(sm_mhz=(s[len(s) // 2] if s else None), sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons),
                    samples=len(s))


# ---------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: oracle/_ref timed on the host cores
def _ref_worker(args):
    scene, n_particles, seed = args
    import oracle_ref
    plan, sc = load_inputs(scene)
    t0 = time.perf_counter()
    r = oracle_ref.mc_run(plan, sc["boxes"], sc["hulls"], n_particles, mt_seed=seed, record=False)
    return time.perf_counter() - t0, r["P"]


def time_reference(scene, n_particles, cores, runs_per_core, seed0=0):
    """`cores` worker processes, each running `runs_per_core` independent runSimulation() calls of the reference
    estimator (single-threaded by construction: MCSimulator.h has no threads) -> rollout-steps/s over the wall
    time of the whole batch."""
    import multiprocessing as mp
    import pcoll_b200
    T = pcoll_b200.load_plan_json(os.path.join(GOLD, "plan_committed.json"))["trajectory"].shape[1]
    jobs = [(scene, n_particles, seed0 + i) for i in range(cores * runs_per_core)]
    t0 = time.perf_counter()
    if cores == 1:
        res = [_ref_worker(j) for j in jobs]
    else:
        with mp.get_context("fork").Pool(cores) as pool:
            res = pool.map(_ref_worker, jobs, chunksize=runs_per_core)
    wall = time.perf_counter() - t0
    steps = len(jobs) * n_particles * T
    return dict(value=steps / wall, wall_s=wall, runs=len(jobs), P_mean=float(np.mean([r[1] for r in res])), T=T)


def reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import oracle_ref
    if not oracle_ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libmcsim_ref.so not built (needs /root/reference at build time)"}))
        return 0
    cores = os.cpu_count() or 1
    time_reference(a.scene, 1000, cores, 1)  # process/page warm-up
    per = []
    for s in range(a.warmup + a.steps):
        r = time_reference(a.scene, a.ref_particles, cores, 4, seed0=1000 * s)
        if s >= a.warmup:
            per.append(r)
    wall = sum(r["wall_s"] for r in per)
    steps = sum(r["runs"] for r in per) * a.ref_particles * per[0]["T"]
    v = steps / wall
    sample = "%d processes x 4 runSimulation() x %d particles x %d waypoints per step (%s scene, mt19937_64 seeds)" % (
        cores, a.ref_particles, per[0]["T"], a.scene)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": 1e3 * wall / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(a, per_gpu=None),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def workload_config(a, per_gpu):
    return {
        "workload": "BASELINE configs[2]: MC convergence run of the committed trajectory.dat/odometry.dat plan (T=56), "
                    "Philox mode, reference estimator law, no early exit; %s rollouts per GPU per step" % (
                        "%d" % per_gpu if per_gpu else "reference arm: bounded sample of the same plan"),
        "scene": a.scene, "waypoints": 56, "rollouts_per_gpu_per_step": per_gpu,
        "collision_query": "every live (hull, box) pair at every step" if getattr(a, "exhaustive", False) else
                           "exact broad phase: identical per-rollout counts to testing every live pair (tests/test_mc_gpu.py::"
                           "test_philox_broad_phase_is_exact); executed pair tests are counted and are what the roofline is charged with",
        "l2": "kernel reads no HBM input (2 KB plan table in shared memory, geometry in __constant__); a 512 MiB buffer "
              "is overwritten between timed steps anyway",
    }


# ---------------------------------------------------------------------------------------------------
def main():
    a = parse()
    if a.impl == "reference":
        return reference_arm(a)

    import torch
    import torch.distributed as dist
    import pcoll_b200
    abi = pcoll_b200.cuda_abi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    plan, sc = load_inputs(a.scene)
    T = plan["trajectory"].shape[1]
    ctx = abi.Context(local)
    ctx.set_scene(sc["boxes"], sc["hulls"])
    stream = torch.cuda.ExternalStream(ctx.stream, device=local)
    n = a.rollouts
    counters = ctx.mc_counters_tensor(T)  # torch view of the device counter block (u64 as int64)
    total = torch.zeros_like(counters)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    fp32_peak, mufu_peak = ctx.measure_peaks(150.0)

    p, keep = abi.make_mc_params(plan, n_rollouts=n, rollout_offset=rank * n, mode=abi.MODE_PHILOX, early_exit=False, exhaustive=a.exhaustive)
    ctx.mc_prepare(p)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step(ev=None):
        with torch.cuda.stream(stream):
            flush.zero_()  # L2 flush, outside the event pair
            if ev:
                ev[0].record(stream)
            ctx.mc_launch(n, rank * n)
            if ev:
                ev[1].record(stream)
            if world > 1:
                dist.all_reduce(counters)  # hits, executed steps, pair tests, first-hit histogram
            total.add_(counters)           # job-wide totals of the timed steps; the block restarts at 0
            counters.zero_()
            if ev:
                ev[2].record(stream)

    for _ in range(a.warmup):
        one_step()
    stream.synchronize()
    total.zero_()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = ctx.launch_count
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(a.steps)]
    t_wall0 = time.perf_counter()
    for k in range(a.steps):
        one_step(evs[k])
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.result()
    launches = ctx.launch_count - launches0
    kernel_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    step_ms = sum(e[0].elapsed_time(e[2]) for e in evs)
    th = total.cpu().numpy().astype(np.uint64)  # after the all-reduce every rank holds the job-wide totals
    tot = dict(hits=int(th[0]), near_contact=int(th[1]), executed_steps=int(th[2]), pair_tests=int(th[3]))
    tmax = torch.tensor([step_ms, kernel_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    step_ms, kernel_ms = float(tmax[0]), float(tmax[1])
    job_steps = tot["executed_steps"]          # whole job, K steps
    assert job_steps == a.steps * world * n * T, (job_steps, a.steps, world, n, T)
    value = job_steps / (step_ms * 1e-3)
    flops_job = job_steps * F_PHILOX_STEP + tot["pair_tests"] * F_SAT
    achieved = flops_job / world / (kernel_ms * 1e-3) / 1e12  # per GPU, TFLOP/s
    P = tot["hits"] / (a.steps * world * n)

    # ---- the same kernel with the broad phase off (every live pair tested at every step, as CheckCollision does):
    # its arithmetic rate under the full SAT load, reported inside `roofline` next to the product configuration
    every_pair = None
    if not a.exhaustive:
        pe, keep_e = abi.make_mc_params(plan, n_rollouts=n, rollout_offset=rank * n, mode=abi.MODE_PHILOX, early_exit=False, exhaustive=True)
        ctx.mc_prepare(pe)
        ev_e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        with torch.cuda.stream(stream):
            ctx.mc_launch(n, rank * n)
            counters.zero_()
            flush.zero_()
            ev_e[0].record(stream)
            ctx.mc_launch(n, rank * n)
            ev_e[1].record(stream)
        stream.synchronize()
        ce = counters.cpu().numpy().astype(np.uint64)
        with torch.cuda.stream(stream):
            counters.zero_()
        stream.synchronize()
        ms_e = ev_e[0].elapsed_time(ev_e[1])
        fl_e = int(ce[2]) * F_PHILOX_STEP + int(ce[3]) * F_SAT
        every_pair = {"rollout_steps_per_s_per_gpu": int(ce[2]) / (ms_e * 1e-3), "pair_tests_per_step": int(ce[3]) / max(int(ce[2]), 1),
                      "achieved": fl_e / (ms_e * 1e-3) / 1e12, "frac": fl_e / (ms_e * 1e-3) / 1e12 / fp32_peak,
                      "hits_equal_product_run": int(ce[0]) * a.steps * world == tot["hits"] if world == 1 else None,
                      "note": "one launch, rank %d, broad phase off (mcsim_mc_params.exhaustive = 1)" % rank}
        ctx.mc_prepare(p)

    # ---- e2e: the plugin command interface with host buffers, every step: parameters + plan pushed as command
    # strings are parsed once; runSimulation uploads the plan table, runs, and returns P as text
    e2e = None
    try:
        e2e = e2e_leg(a, plan, sc, local, rank, world, barrier)
    except Exception as exc:  # the leg is reported, never silently dropped
        e2e = {"value": None, "unit": UNIT, "error": repr(exc)}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": step_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(a, n),
        "P_collision": P, "wall_ms_per_step_incl_flush": 1e3 * t_wall / a.steps,
        "clocks": clocks, "gpu_launches": launches, "e2e": e2e,
        "roofline": {
            "bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
            "traffic": None,
            "traffic_note": "no HBM input: tables in __constant__ / shared memory, state in registers; ncu capture of a 4e6-rollout "
                            "launch: dram__bytes_read 52 KB, dram__bytes_write 0 (profiles/r01_philox2_ladder_ncu_summary.csv)",
            "kernel": "mc_philox2_kernel<no early exit, reference estimator, L=8> (two rollouts per thread, packed FP32 pairs)",
            "kernel_ms_per_launch": kernel_ms / a.steps,
            "algorithmic": "%d flops per Philox rollout-step + %d per SAT pair test x %.1f live pair tests per step" % (
                F_PHILOX_STEP, F_SAT, tot["pair_tests"] / job_steps),
            "sfu": {"achieved": job_steps * SFU_PHILOX_STEP / world / (kernel_ms * 1e-3) / 1e12, "peak": mufu_peak, "unit": "Top/s",
                    "frac": job_steps * SFU_PHILOX_STEP / world / (kernel_ms * 1e-3) / 1e12 / mufu_peak,
                    "algorithmic": "%.1f MUFU operations per Philox rollout-step; the FP32 and SFU pipes run side by side, the fractions do not add" % SFU_PHILOX_STEP},
            "peak_source": "FFMA-chain microbenchmark (csrc/peaks_sm100.cuh) measured in this process; MUFU peak %.3f Top/s" % mufu_peak,
            "file_peaks": read_measured_peaks(),
            "every_pair_tested": every_pair,
        },
    }
    if rank == 0 and not a.no_extras:
        line["extras"] = extras(a, ctx, abi, plan, T, stream, torch)
        try:
            line["extras"].update(extras_scoring_em(ctx, abi, plan, sc, fp32_peak))
        except Exception as exc:  # reported, never silently dropped
            line["extras"]["scoring_em_error"] = repr(exc)
    if rank == 0 and "extras" in line and line["extras"].get("gmm_batch_main") in line["extras"]:
        key = line["extras"]["gmm_batch_main"]
        g = line["extras"][key]
        line["secondary"] = {"metric": "gmm_pcoll_trajectories_per_s", "value": g["trajectories_per_s"], "unit": "trajectories/s",
                             "config": "runGMMEstimation, K = 3, 10^4 samples per waypoint, 56 waypoints, %s runs per batch "
                                       "(mcsim_cuda_gmm_run_batch), host buffers in and out (BASELINE configs[1])" % key.split("batch")[-1],
                             "kernel_only": g["trajectories_per_s_kernel_only"],
                             "single_run_value": line["extras"].get("gmm_K3_N10000", {}).get("trajectories_per_s")}
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        import oracle_ref
        if oracle_ref.available():
            r = time_reference(a.scene, a.ref_particles, 1, 24)
            line["cpu_baseline"] = {
                "value": r["value"], "unit": UNIT, "cores": 1, "kind": "reference",
                "sample": "24 x runSimulation() of the reference estimator headers (oracle/_ref) x %d particles x %d "
                          "waypoints, 1 thread (the reference is single-threaded), CPU SAT on the same geometry; %.1f s"
                          % (a.ref_particles, T, r["wall_s"]),
                "P_mean": r["P_mean"]}
            # the GMM mode of the reference on the same core: runGMMEstimation, K = 3, 10^4 samples per waypoint
            plan_c, sc_c = load_inputs(a.scene)
            t0 = time.perf_counter()
            gp = [oracle_ref.gmm_run(plan_c, sc_c["boxes"], sc_c["hulls"], 3, 10000, mt_seed=s_, multinomial_seed=s_ + 1, record=False)["P"] for s_ in range(8)]
            dt = time.perf_counter() - t0
            line["cpu_baseline"]["gmm"] = {"value": 8 / dt, "unit": "trajectories/s", "cores": 1, "kind": "reference",
                                           "sample": "8 x runGMMEstimation() of the reference estimator headers, K = 3, 10^4 samples x 56 waypoints, 1 thread; %.1f s" % dt,
                                           "P_mean": float(np.mean(gp))}
            # the vendored Armadillo EM (gmm_full::learn, keep_existing) on one core: K = 3, 2 x 10^5 samples, 10 iterations
            xe, he, me, ce = pcoll_b200.inputs.synthetic_em_problem(9000, 200000, 3)
            t0 = time.perf_counter()
            oracle_ref.em_fit(xe.T.astype(np.float64), me, ce.reshape(3, 9), he, 10)
            dt = time.perf_counter() - t0
            line["cpu_baseline"]["em"] = {"value": 200000 * 10 / dt, "unit": "sample-iterations/s", "cores": 1, "kind": "reference",
                                          "sample": "arma::gmm_full::learn(keep_existing, em_iter = 10), K = 3, 2 x 10^5 samples, 1 thread; %.1f s" % dt}
        else:
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    if rank == 0:
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def read_measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return {k: d.get(k) for k in ("hbm_gbs", "bf16_tflops", "sm_max_mhz")}
    except Exception:
        return None


def e2e_leg(a, plan, sc, local, rank, world, barrier):
    """rollout-steps/s through the reference-facing command interface (MCModule.SendCommand), host buffers."""
    import torch
    import pcoll_b200
    T = plan["trajectory"].shape[1]
    n = a.e2e_rollouts
    mod = pcoll_b200.MCModule(scene=sc, device=local)
    pcoll_b200.push_parameters(mod, plan, num_particles=n)
    mod.SendCommand("setMode philox")
    mod.SendCommand("setEarlyExit 0")
    mod.SendCommand("setRolloutOffset %d" % (rank * n))
    for _ in range(max(1, a.warmup)):
        mod.SendCommand("runSimulation")
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        s = mod.SendCommand("runSimulation")
        float(s)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    io = mod.io_bytes()
    mod.close()
    return {"value": a.steps * world * n * T / float(t[0]), "unit": UNIT, "h2d_bytes_per_step": io[0], "d2h_bytes_per_step": io[1],
            "rollouts_per_gpu_per_step": n, "api": "MCModule.SendCommand('runSimulation') -> libmcsim.so -> libmcsim_cuda.so",
            "timing": "host wall clock around K synchronous calls, max over ranks"}


def extras(a, ctx, abi, plan, T, stream, torch):
    """Side measurements on rank 0 (not the headline): early-exit MC, the GMM mode (trajectories/s)."""
    out = {}
    n = min(a.rollouts, 10 ** 7)
    p, keep = abi.make_mc_params(plan, n_rollouts=n, mode=abi.MODE_PHILOX, early_exit=True)
    ctx.mc_run(p)
    r = ctx.mc_run(p)
    out["mc_early_exit"] = {"nominal_rollout_steps_per_s": n * T / (r["kernel_ms"] * 1e-3),
                            "executed_rollout_steps_per_s": r["executed_steps"] / (r["kernel_ms"] * 1e-3),
                            "rollouts": n, "P": r["P"]}
    if hasattr(ctx, "gmm_run"):
        try:
            for K, ns in ((3, 10000), (3, 1000000)):
                ctx.gmm_run(plan, K, ns, mode=abi.MODE_PHILOX)
                t0 = time.perf_counter()
                reps = 5
                for i in range(reps):
                    g = ctx.gmm_run(plan, K, ns, mode=abi.MODE_PHILOX, seed=1 + i)
                dt = (time.perf_counter() - t0) / reps
                out["gmm_K%d_N%d" % (K, ns)] = {"trajectories_per_s": 1.0 / dt, "sample_checks_per_s": T * ns / dt,
                                                "kernel_ms": g["kernel_ms"], "P": g["P"]}
            # batches: 256 runs, 1,024 runs, and 9 x the SM count (three full waves of the one-CTA-per-run kernel at
            # its three resident CTAs per SM: 1,332 on a 148-SM B200) -- the last is the `secondary` metric
            n_sm = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
            for ns, B in ((10000, 256), (10000, 1024), (10000, 9 * n_sm)):
                ctx.gmm_run_batch(plan, 3, ns, B)
                best = None
                for rep in range(3):
                    t0 = time.perf_counter()
                    g = ctx.gmm_run_batch(plan, 3, ns, B, stream0=B * (rep + 1))
                    dt = time.perf_counter() - t0
                    best = dt if best is None or dt < best else best
                dt = best
                out["gmm_K3_N%d_batch%d" % (ns, B)] = {
                    "trajectories_per_s": B / dt, "trajectories_per_s_kernel_only": B / (g["kernel_ms"] * 1e-3),
                    "sample_checks_per_s": B * T * ns / dt, "kernel_ms": g["kernel_ms"],
                    "pair_tests_per_sample": g["pair_tests"] / (B * T * ns),
                    "P_mean": float(g["P"].mean()), "P_std": float(g["P"].std()), "timing": "best of 3 calls, host wall clock"}
                out["gmm_batch_main"] = "gmm_K3_N%d_batch%d" % (ns, B)
        except Exception as exc:
            out["gmm_error"] = repr(exc)
    return out


def extras_scoring_em(ctx, abi, plan, sc, fp32_peak):
    """BASELINE configs 4 and 5 as side measurements on rank 0 (subsets stated in each record)."""
    import pcoll_b200
    inp = pcoll_b200.inputs
    out = {}
    peaks = read_measured_peaks() or {}
    # ---- config 4: batched plan scoring, 128 of the 4,096 plans x 65,536 rollouts x 500 waypoints, 256 boxes
    boxes = inp.synthetic_obstacle_field(4096, 256)
    n_plans, T4, R = 128, 500, 65536
    tr, od = inp.synthetic_plans(4096, n_plans, T4, boxes)
    ctx.set_scene(boxes, sc["hulls"])
    for name, margin, plans, rollouts in (("score_config4_broad_phase", 3.0, n_plans, R), ("score_config4_exhaustive", -1.0, 16, 16384)):
        ctx.score_prepare(plan, tr[:plans], od[:plans], rollouts, margin=margin)
        ctx.score_launch()
        ctx.score_collect()
        ev = [cuda_event(ctx) for _ in range(2)]
        ev[0].record()
        ctx.score_launch()
        ev[1].record()
        r = ctx.score_collect()
        ms = ev[0].elapsed_ms(ev[1])
        flops = r["executed_steps"] * F_PHILOX_STEP + r["pair_tests"] * F_SAT
        out[name] = {
            "subset": "%d plans x %d rollouts x %d waypoints, 256-box field, margin %s" % (plans, rollouts, T4, margin),
            "nominal_rollout_steps_per_s": plans * rollouts * T4 / (ms * 1e-3), "kernel_ms": ms,
            "pair_tests_per_rollout_step": r["pair_tests"] / max(r["executed_steps"], 1),
            "algorithmic_tflops": flops / (ms * 1e-3) / 1e12, "frac_of_fp32_peak": flops / (ms * 1e-3) / 1e12 / fp32_peak,
            "P_mean": float(r["P"].mean()), "P_max": float(r["P"].max())}
    ctx.set_scene(sc["boxes"], sc["hulls"])
    # ---- config 5: EM sweep, 56 waypoint buffers x 10^6 FP32 samples (672 MB resident), 10 iterations
    F, n, iters = 56, 10 ** 6, 10
    traj = plan["trajectory"]
    base = [inp.synthetic_em_problem(9000 + f, n, 16, centre=(0.0, 0.0, 0.0)) for f in range(8)]
    probs = []  # 56 distinct buffers: 8 seeded mixtures, each moved to 7 different trajectory poses
    for f in range(F):
        b, ctr = base[f % 8], traj[:, f % traj.shape[1]]
        probs.append(((b[0] + ctr.astype(np.float32)[:, None]).astype(np.float32), b[1], b[2] + ctr[None, :], b[3]))
    x = np.stack([q[0] for q in probs])
    first = True
    for K in (1, 2, 3, 4, 8, 16):
        h = np.stack([np.full(K, 1.0 / K) for q in probs]); m = np.stack([q[2][:K] for q in probs]); c = np.stack([q[3][:K] for q in probs])
        ctx.em_fit(x if first else None, h, m, c, 2)
        first = False
        r = ctx.em_fit(None, h, m, c, iters)
        sec = r["kernel_ms"] * 1e-3
        gbs = 12.0 * n * F * iters / sec / 1e9
        out["em_config5_K%d" % K] = {
            "fits": F, "samples_per_fit": n, "iterations": iters, "kernel_ms": r["kernel_ms"],
            "sample_iterations_per_s": n * F * iters / sec, "algorithmic_GBps": gbs,
            "frac_of_hbm_peak": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
            "algorithmic_tflops": 61.0 * K * n * F * iters / sec / 1e12,
            "frac_of_fp32_peak": 61.0 * K * n * F * iters / sec / 1e12 / fp32_peak,
            "avg_log_p_last": float(np.nanmean(r["avg_log_p"][:, -1]))}
    return out


class cuda_event:
    """CUDA event on the library's own stream (torch.cuda.Event sees only torch's current stream)."""
    def __init__(self, ctx):
        import torch
        self._torch = torch
        self._stream = torch.cuda.ExternalStream(ctx.stream, device=ctx.device)
        self._ev = torch.cuda.Event(enable_timing=True)

    def record(self):
        self._ev.record(self._stream)

    def elapsed_ms(self, other):
        other._ev.synchronize()
        return self._ev.elapsed_time(other._ev)


if __name__ == "__main__":
    sys.exit(main())
