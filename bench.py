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
`--rollouts` rollouts IN TOTAL (`--scaling strong`, the default: configs[2] is a fixed job of 10^8 rollouts "sharded across
1/2/4/8 B200"; rank r owns rollout ids [r*N/G, (r+1)*N/G), Philox counters come from the global id, so hits and
P_collision are the same at every N) or PER GPU (`--scaling weak`), followed by the NCCL all-reduce of the hit
counters on the same stream.

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
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong: --rollouts is the whole job, split over the GPUs (BASELINE configs[2]); weak: --rollouts per GPU")
    ap.add_argument("--rollouts", type=int, default=10 ** 8, help="rollouts per step: of the whole job (strong) or per GPU (weak)")
    ap.add_argument("--e2e-rollouts", type=int, default=8 * 10 ** 7, help="rollouts per e2e step: whole job (strong) or x GPUs (weak: this / 8 per GPU)")
    ap.add_argument("--group-e2e-child", type=int, default=0, help=argparse.SUPPRESS)
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


def workload_config(a, per_gpu, job=None):
    return {
        "workload": "BASELINE configs[2]: MC convergence run of the committed trajectory.dat/odometry.dat plan (T=56), "
                    "Philox mode, reference estimator law, no early exit; %s" % (
                        ("%d rollouts per step in total, %s scaling (%d per GPU)" % (job, a.scaling, per_gpu)) if per_gpu
                        else "reference arm: bounded sample of the same plan"),
        "scene": a.scene, "waypoints": 56, "rollouts_per_gpu_per_step": per_gpu, "rollouts_per_step": job,
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
    if a.group_e2e_child:
        return group_e2e_child(a)

    import torch
    import torch.distributed as dist
    import pcoll_b200
    abi = pcoll_b200.cuda_abi
    from pcoll_b200.sharding import shard_range

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
    # strong scaling: one fixed job, rollout ids [lo, hi) per rank; weak: --rollouts per rank
    n_job = a.rollouts if a.scaling == "strong" else a.rollouts * world
    lo, hi = shard_range(n_job, rank, world)
    n = hi - lo
    counters = ctx.mc_counters_tensor(T)  # torch view of the device counter block (u64 as int64)
    total = torch.zeros_like(counters)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    fp32_peak, mufu_peak = ctx.measure_peaks(150.0)

    p, keep = abi.make_mc_params(plan, n_rollouts=n, rollout_offset=lo, mode=abi.MODE_PHILOX, early_exit=False, exhaustive=a.exhaustive)
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
            ctx.mc_launch(n, lo)
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
    assert job_steps == a.steps * n_job * T, (job_steps, a.steps, n_job, T)
    value = job_steps / (step_ms * 1e-3)
    flops_job = job_steps * F_PHILOX_STEP + tot["pair_tests"] * F_SAT
    achieved = flops_job / world / (kernel_ms * 1e-3) / 1e12  # per GPU, TFLOP/s
    assert tot["hits"] % a.steps == 0           # every step is the same job: the same hits
    hits_per_step = tot["hits"] // a.steps
    P = hits_per_step / n_job

    # ---- BASELINE configs[2]'s invariant, checked inside the bench: the sharded job's hit total and first-hit histogram equal
    # one GPU's on the same rollout ids (a 2^22-rollout job, so that rank 0 can run all of it alone in milliseconds)
    invariance = None
    if world > 1:
        n_small = 1 << 22
        slo, shi = shard_range(n_small, rank, world)
        ps, keep_s = abi.make_mc_params(plan, n_rollouts=shi - slo, rollout_offset=slo, mode=abi.MODE_PHILOX, early_exit=False)
        ctx.mc_prepare(ps)
        with torch.cuda.stream(stream):
            counters.zero_()
            ctx.mc_launch(shi - slo, slo)
            dist.all_reduce(counters)
        stream.synchronize()
        sharded = counters.cpu().numpy().astype(np.uint64)
        with torch.cuda.stream(stream):
            counters.zero_()
        if rank == 0:
            pa, keep_a = abi.make_mc_params(plan, n_rollouts=n_small, rollout_offset=0, mode=abi.MODE_PHILOX, early_exit=False)
            alone = ctx.mc_run(pa, want_hist=True)
            same = int(sharded[0]) == alone["hits"] and np.array_equal(sharded[4:4 + T], alone["first_hit_hist"])
            invariance = {"rollouts": n_small, "hits_sharded": int(sharded[0]), "hits_one_gpu": alone["hits"], "histogram_equal": bool(same)}
            assert same, invariance
        ctx.mc_prepare(p)
        barrier()

    # ---- the same kernel with the broad phase off (every live pair tested at every step, as CheckCollision does):
    # its arithmetic rate under the full SAT load, reported inside `roofline` next to the product configuration
    every_pair = None
    if not a.exhaustive:
        ne = min(n, 25 * 10 ** 6)
        pe, keep_e = abi.make_mc_params(plan, n_rollouts=ne, rollout_offset=lo, mode=abi.MODE_PHILOX, early_exit=False, exhaustive=True)
        ctx.mc_prepare(pe)
        ev_e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        with torch.cuda.stream(stream):
            ctx.mc_launch(ne, lo)
            counters.zero_()
            flush.zero_()
            ev_e[0].record(stream)
            ctx.mc_launch(ne, lo)
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
                      "note": "one launch of %d rollouts, rank %d, broad phase off (mcsim_mc_params.exhaustive = 1)" % (ne, rank)}
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
        "ms_per_step": step_ms / a.steps, "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(a, n, n_job),
        "P_collision": P, "hits_per_step": hits_per_step, "hits_invariance": invariance,
        "wall_ms_per_step_incl_flush": 1e3 * t_wall / a.steps,
        "clocks": clocks, "gpu_launches": launches, "e2e": e2e,
        "roofline": {
            "bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak,
            "traffic": None,
            "traffic_note": "no HBM input: tables in __constant__ / shared memory, state in registers; ncu capture of a 4e6-rollout "
                            "launch: dram__bytes_read 52 KB, dram__bytes_write 0 (profiles/r01_philox2_ladder_ncu_summary.csv)",
            "kernel": "mc_philox2_kernel<no early exit, reference estimator, L=8> (two rollouts per thread, packed FP32 pairs)",
            "kernel_ms_per_launch": kernel_ms / a.steps,
            "algorithmic": "%d flops per Philox rollout-step + %d per SAT pair test x %.2f executed pair tests per step" % (
                F_PHILOX_STEP, F_SAT, tot["pair_tests"] / job_steps),
            "sfu": {"achieved": job_steps * SFU_PHILOX_STEP / world / (kernel_ms * 1e-3) / 1e12, "peak": mufu_peak, "unit": "Top/s",
                    "frac": job_steps * SFU_PHILOX_STEP / world / (kernel_ms * 1e-3) / 1e12 / mufu_peak,
                    "algorithmic": "%.1f MUFU operations per Philox rollout-step; the FP32 and SFU pipes run side by side, the fractions do not add" % SFU_PHILOX_STEP},
            "peak_source": "FFMA-chain microbenchmark (csrc/peaks_sm100.cuh) measured in this process; MUFU peak %.3f Top/s" % mufu_peak,
            "file_peaks": read_measured_peaks(),
            "every_pair_tested": every_pair,
        },
    }
    if not a.no_extras:
        # legs every rank takes part in (GMM replicas, sharded EM); rank 0 keeps the records
        try:
            multi = extras_all_ranks(a, ctx, abi, plan, T, torch, dist, world, rank, barrier, fp32_peak)
        except Exception as exc:
            multi = {"all_ranks_error": repr(exc)}
        if rank == 0:
            line["extras"] = extras(a, ctx, abi, plan, T, stream, torch)
            line["extras"].update(multi)
            try:
                line["extras"].update(extras_scoring_em(ctx, abi, plan, sc, fp32_peak))
            except Exception as exc:  # reported, never silently dropped
                line["extras"]["scoring_em_error"] = repr(exc)
            try:
                line["extras"]["config1_shared_mode"] = config1_leg(a, plan, sc, local)
            except Exception as exc:
                line["extras"]["config1_error"] = repr(exc)
    if rank == 0 and "extras" in line and "gmm_replicas" in line["extras"]:
        g = line["extras"]["gmm_replicas"]
        line["secondary"] = {"metric": "gmm_pcoll_trajectories_per_s", "value": g["trajectories_per_s"], "unit": "trajectories/s", "n_gpus": world,
                             "config": "runGMMEstimation, K = 3, 10^4 samples per waypoint, 56 waypoints, %d runs per batch and GPU "
                                       "(mcsim_cuda_gmm_run_batch; replicas: every GPU its own batch, no collective), host buffers in and out "
                                       "(BASELINE configs[1])" % g["runs_per_gpu"],
                             "kernel_only": g["trajectories_per_s_kernel_only"],
                             "single_run_value": line["extras"].get("gmm_K3_N10000", {}).get("trajectories_per_s"),
                             "roofline": g["roofline"]}
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
    ctx.close()
    del counters, total, flush
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
        if rank != 0:
            return 0
        # ---- e2e at N GPUs through the C ABI alone: ONE plugin module, `setDevices 0 .. N-1` (NCCL communicators owned
        # by the handle), driven from a torch-free child process once the other ranks have left their GPUs
        torch.cuda.synchronize()
        line["e2e"] = group_e2e_parent(a, world, line["e2e"])
    if rank == 0:
        print(json.dumps(line))
    return 0


def group_e2e_parent(a, world, per_rank):
    import subprocess as sp
    time.sleep(1.0)
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "GROUP_RANK",
                                                            "ROLE_RANK", "ROLE_WORLD_SIZE", "LOCAL_WORLD_SIZE", "TORCHELASTIC_RUN_ID")}
    cmd = [sys.executable, os.path.abspath(__file__), "--group-e2e-child", str(world), "--steps", str(a.steps), "--warmup", str(a.warmup),
           "--scene", a.scene, "--e2e-rollouts", str(a.e2e_rollouts), "--scaling", a.scaling]
    try:
        out = sp.run(cmd, env=env, capture_output=True, text=True, timeout=240)
        rec = json.loads(out.stdout.strip().splitlines()[-1])
        rec["per_rank_modules"] = per_rank    # the same job as N processes x one module each (torch.distributed around them)
        return rec
    except Exception as exc:
        per_rank = dict(per_rank or {})
        per_rank["single_process_group_error"] = repr(exc)
        return per_rank


def em_group_leg(G, n=1_000_000, iters=10, reps=5):
    """Config 5's fit sharded over the G devices of ONE handle (mcsim_cuda_group_em_fit, samples resident): microseconds per
    iteration with the statistics exchange fused into the M-step kernel (peer-memory push + gather) and with one
    ncclAllReduce per iteration.  device = max over the devices of the CUDA-event time of the E / M launches; wall = host
    clock around the call (parameter upload, download and, for the fused path, the per-device host threads included)."""
    import pcoll_b200
    abi, inp = pcoll_b200.cuda_abi, pcoll_b200.inputs
    out = {}
    try:
        grp = abi.Group(list(range(G)))
    except Exception as exc:
        return {"error": repr(exc)}
    try:
        out["n"], out["iters"], out["default_exchange"] = n, iters, "peer" if grp.em_exchange == abi.EM_EXCHANGE_PEER else "nccl"
        for K in (1, 3, 16):
            x, h, m, c = inp.synthetic_em_problem(9000 + K, n, K)
            rec = {}
            for name, mode in (("peer", abi.EM_EXCHANGE_PEER), ("nccl", abi.EM_EXCHANGE_NCCL)):
                try:
                    grp.set_em_exchange(mode)
                except abi.McsimCudaError as exc:
                    rec[name] = {"unavailable": str(exc)}
                    continue
                grp.em_fit(x, h, m, c, iters)
                dev, wall = [], []
                for _ in range(reps):
                    t0 = time.perf_counter()
                    r = grp.em_fit(None, h, m, c, iters)
                    wall.append((time.perf_counter() - t0) * 1e6 / iters)
                    dev.append(r["kernel_ms"] * 1e3 / iters)
                rec[name] = {"device_us_per_iter": float(np.median(dev)), "wall_us_per_iter": float(np.median(wall)),
                             "iters_run": int(r["iters_run"][0]), "heft0": float(r["hefts"][0, 0])}
            out["K%d" % K] = rec
    finally:
        grp.close()
    return out


def group_e2e_child(a):
    """Torch-free: one MCModule over `--group-e2e-child` devices; runSimulation of the whole e2e job, host wall clock."""
    import pcoll_b200
    G = a.group_e2e_child
    plan, sc = load_inputs(a.scene)
    T = plan["trajectory"].shape[1]
    n_job = a.e2e_rollouts if a.scaling == "strong" else (a.e2e_rollouts // 8) * G
    mod = pcoll_b200.MCModule(scene=sc, device=0)
    mod.SendCommand("setDevices " + " ".join(str(d) for d in range(G)))
    pcoll_b200.push_parameters(mod, plan, num_particles=n_job)
    mod.SendCommand("setMode philox")
    mod.SendCommand("setEarlyExit 0")
    mod.SendCommand("setSeed %d" % 0x5AFEC0DE)
    for _ in range(max(1, a.warmup)):
        mod.SendCommand("runSimulation")
    t0 = time.perf_counter()
    for _ in range(a.steps):
        float(mod.SendCommand("runSimulation"))
    dt = time.perf_counter() - t0
    io = mod.io_bytes()
    hits = int(mod.SendCommand("getStats").split()[0])
    mod.close()
    print(json.dumps({"value": a.steps * n_job * T / dt, "unit": UNIT, "h2d_bytes_per_step": io[0] * G, "d2h_bytes_per_step": io[1] * G,
                      "rollouts_per_step": n_job, "n_gpus": G, "hits_last_step": hits, "em_group": em_group_leg(G),
                      "api": "ONE MCModule, SendCommand('setDevices 0..%d') + SendCommand('runSimulation') -> libmcsim.so -> mcsim_cuda_group_mc_run "
                             "(rollout ids split over the devices, ncclAllReduce of the counters on the handle's communicators); no torch in this process" % (G - 1),
                      "timing": "host wall clock around K synchronous calls"}))
    return 0


def config1_leg(a, plan, sc, local):
    """BASELINE configs[0] as the reference driver runs it: runSimulation in the plugin's default `shared` mode (one
    closed-loop real robot per run, MCSimulator.h:754-761) at N = 1,000 (BASELINE) and 10,000 (MCSimulation.py:164):
    microseconds per call through SendCommand, beside the reference estimator on one host core."""
    import pcoll_b200
    out = {}
    for N in (1000, 10000):
        mod = pcoll_b200.MCModule(scene=sc, device=local)
        pcoll_b200.push_parameters(mod, plan, num_particles=N)
        mod.SendCommand("setSeed 1")
        for _ in range(20):
            mod.SendCommand("runSimulation")
        reps = 200
        t0 = time.perf_counter()
        Ps = [float(mod.SendCommand("runSimulation")) for _ in range(reps)]
        dt = (time.perf_counter() - t0) / reps
        kernel_ms = float(mod.SendCommand("getStats").split()[4])
        mod.close()
        rec = {"us_per_call": dt * 1e6, "kernel_us": kernel_ms * 1e3, "P_mean_of_200_runs": float(np.mean(Ps)), "P_std": float(np.std(Ps, ddof=1)),
               "rollout_steps_per_s": N * plan["trajectory"].shape[1] / dt}
        try:
            import oracle_ref
            if oracle_ref.available():
                t0 = time.perf_counter()
                k = 3 if N > 1000 else 10
                for s_ in range(k):
                    oracle_ref.mc_run(plan, sc["boxes"], sc["hulls"], N, mt_seed=s_, record=False)
                rec["reference_us_per_call_one_core"] = (time.perf_counter() - t0) / k * 1e6
        except Exception:
            pass
        out["N%d" % N] = rec
    return out


def extras_all_ranks(a, ctx, abi, plan, T, torch, dist, world, rank, barrier, fp32_peak):
    """The legs of BASELINE.json's metric that need every rank: GMM trajectories/s as replicas (each GPU its own batch of
    runs, no collective -- one run is a chain of 56 dependent waypoints and does not shard) and the EM fit
    with its samples sharded (one all-reduce of the K 10 + 1 statistics per fit and iteration)."""
    from pcoll_b200.sharding import em_fit_sharded, shard_range
    import pcoll_b200
    out = {}
    n_sm = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
    B, ns = 8 * n_sm, 10000   # two full waves of the 128-thread shape (4 runs in flight per SM)
    ctx.gmm_run_batch(plan, 3, ns, B, stream0=rank * 100000)
    best, rec = None, None
    for rep in range(3):
        barrier()
        t0 = time.perf_counter()
        g = ctx.gmm_run_batch(plan, 3, ns, B, stream0=rank * 100000 + B * (rep + 1))
        dt = time.perf_counter() - t0
        t = torch.tensor([dt, g["kernel_ms"] * 1e-3], dtype=torch.float64, device="cuda")
        tests = torch.tensor([float(g["pair_tests"])], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(tests)
        if best is None or float(t[0]) < best:
            best, rec = float(t[0]), (float(t[1]), float(tests[0]), g)
    ksec, tests, g = rec
    checks = world * B * T * ns
    flops = checks * F_GMM_SAMPLE + tests * F_SAT
    out["gmm_replicas"] = {
        "trajectories_per_s": world * B / best, "trajectories_per_s_kernel_only": world * B / ksec, "runs_per_gpu": B, "n_gpus": world,
        "sample_checks_per_s": checks / best, "pair_tests_per_sample": tests / checks, "P_mean": float(g["P"].mean()), "P_std": float(g["P"].std()),
        "timing": "best of 3 barrier-aligned calls, host wall clock, max over ranks",
        "roofline": {"bound": "fp32", "achieved": flops / world / ksec / 1e12, "peak": fp32_peak, "unit": "TFLOP/s", "frac": flops / world / ksec / 1e12 / fp32_peak,
                     "kernel": "gmm_member_run_kernel (one CTA per run for the whole plan)",
                     "algorithmic": "%d flops per sample (draw, L z, pose, moment sums) + %d per executed SAT pair test x %.2f tests per sample" % (
                         F_GMM_SAMPLE, F_SAT, tests / checks)}}
    # ---- EM, samples sharded: BASELINE config 5's 56 waypoint buffers x 10^6 samples, K = 1, 3, 16, 10 iterations
    F, n_tot, iters = 56, 10 ** 6, 10
    lo, hi = shard_range(n_tot, rank, world)
    lo, hi = lo & ~3, (hi & ~3) if rank + 1 < world else n_tot
    xf, m16, c16 = config5_buffers(plan, n_tot, F)
    x = np.ascontiguousarray(xf[:, :, lo:hi])
    del xf
    first = True
    for K in (1, 3, 16):
        h = np.full((F, K), 1.0 / K); m = np.ascontiguousarray(m16[:, :K]); c = np.ascontiguousarray(c16[:, :K])
        em_fit_sharded(ctx, x if first else None, n_tot, h, m, c, 2)
        first = False
        times = []
        for with_collective in (True, False):   # the same loop with and without the all-reduce: its share of the iteration
            ctx.em_prepare(None, h, m, c, iters, n_total=n_tot)
            barrier()
            evs = [cuda_event(ctx) for _ in range(2)]
            evs[0].record()
            for _ in range(iters):
                ctx.em_estep()
                if with_collective and world > 1:
                    with torch.cuda.stream(ctx.torch_stream):
                        dist.all_reduce(ctx.em_stats_tensor())
                ctx.em_mstep()
            evs[1].record()
            times.append(evs[0].elapsed_ms(evs[1]))
            if with_collective:
                r = ctx.em_collect()
        ms, ms_local = times
        t = torch.tensor([ms, ms_local], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_local = float(t[0]), float(t[1])
        out["em_sharded_K%d" % K] = {
            "fits": F, "samples_per_fit": n_tot, "iterations": iters, "n_gpus": world, "ms": ms, "ms_without_all_reduce": ms_local,
            "all_reduce_share": max(0.0, 1.0 - ms_local / ms) if ms > 0 else None,
            "sample_iterations_per_s": F * n_tot * iters / (ms * 1e-3), "iters_run_min_max": [int(r["iters_run"].min()), int(r["iters_run"].max())],
            "collective": "all_reduce(sum, float64) of %d doubles per iteration (NCCL, on the library's stream)" % (F * (K * 10 + 1))}
    return out


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
    from pcoll_b200.sharding import shard_range
    T = plan["trajectory"].shape[1]
    n_job = a.e2e_rollouts if a.scaling == "strong" else (a.e2e_rollouts // 8) * world
    lo, hi = shard_range(n_job, rank, world)
    n = hi - lo
    mod = pcoll_b200.MCModule(scene=sc, device=local)
    pcoll_b200.push_parameters(mod, plan, num_particles=n)
    mod.SendCommand("setMode philox")
    mod.SendCommand("setEarlyExit 0")
    mod.SendCommand("setSeed %d" % 0x5AFEC0DE)
    mod.SendCommand("setRolloutOffset %d" % lo)
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
    return {"value": a.steps * n_job * T / float(t[0]), "unit": UNIT, "h2d_bytes_per_step": io[0] * world, "d2h_bytes_per_step": io[1] * world,
            "rollouts_per_step": n_job, "rollouts_per_gpu_per_step": n, "api": "MCModule.SendCommand('runSimulation') -> libmcsim.so -> libmcsim_cuda.so",
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
                            "executed_fraction": r["executed_steps"] / (n * T), "rollouts": n, "P": r["P"], "hits": r["hits"],
                            "how": "survivor compaction: one launch per 4 waypoints, the rollouts still free are compacted in device memory "
                                   "(mc_philox2_segment_kernel); hits identical to the no-exit run"}
    pn, keep_n = abi.make_mc_params(plan, n_rollouts=n, mode=abi.MODE_PHILOX, early_exit=True, no_compaction=True)
    ctx.mc_run(pn)
    rn = ctx.mc_run(pn)
    out["mc_early_exit"]["one_launch_nominal_rollout_steps_per_s"] = n * T / (rn["kernel_ms"] * 1e-3)
    assert rn["hits"] == r["hits"]
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
            for ns, B in ((10000, 256), (10000, 1024)):
                ctx.gmm_run_batch(plan, 3, ns, B)
                best = None
                for rep in range(3):
                    t0 = time.perf_counter()
                    g = ctx.gmm_run_batch(plan, 3, ns, B, stream0=B * (rep + 1))
                    dt = time.perf_counter() - t0
                    best = dt if best is None or dt < best else best
                out["gmm_K3_N%d_batch%d" % (ns, B)] = {
                    "trajectories_per_s": B / best, "trajectories_per_s_kernel_only": B / (g["kernel_ms"] * 1e-3), "kernel_ms": g["kernel_ms"],
                    "pair_tests_per_sample": g["pair_tests"] / (B * T * ns), "timing": "best of 3 calls, host wall clock"}
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
    x, m16, c16 = config5_buffers(plan, n, F)
    first = True
    for K in (1, 2, 3, 4, 8, 16):
        h = np.full((F, K), 1.0 / K); m = np.ascontiguousarray(m16[:, :K]); c = np.ascontiguousarray(c16[:, :K])
        ctx.em_fit(x if first else None, h, m, c, 2)
        first = False
        r = ctx.em_fit(None, h, m, c, iters)
        sec = r["kernel_ms"] * 1e-3
        done_it = int(r["iters_run"].astype(np.int64).sum())      # executed fit-iterations (= F * iters unless a fit converged early)
        gbs = 12.0 * n * done_it / sec / 1e9
        out["em_config5_K%d" % K] = {
            "fits": F, "samples_per_fit": n, "iterations": iters, "kernel_ms": r["kernel_ms"],
            "sample_iterations_per_s": n * done_it / sec, "algorithmic_GBps": gbs,
            "frac_of_hbm_peak": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
            "algorithmic_tflops": 61.0 * K * n * done_it / sec / 1e12,
            "frac_of_fp32_peak": 61.0 * K * n * done_it / sec / 1e12 / fp32_peak,
            "iters_run_min_max": [int(r["iters_run"].min()), int(r["iters_run"].max())],
            "executed_sample_iterations": int(r["iters_run"].astype(np.int64).sum()) * n,
            "note": "rates are charged with the EXECUTED sample-iterations (a fit that converges stops; iters_run says whether any did)",
            "avg_log_p_last": float(np.nanmean(r["avg_log_p"][:, -1]))}
    return out


def config5_buffers(plan, n=10 ** 6, F=56):
    """BASELINE config 5: F distinct waypoint buffers of n FP32 samples (672 MB at 56 x 10^6, so an iteration streams from
    HBM, not L2): 8 seeded 3-component mixtures, each moved to 7 trajectory poses.  Returns x [F, 3, n] and the
    keep_existing starting points for up to 16 components."""
    import pcoll_b200
    traj = plan["trajectory"]
    base = [pcoll_b200.inputs.synthetic_em_problem(9000 + f, n, 16, centre=(0.0, 0.0, 0.0)) for f in range(8)]
    x = np.empty((F, 3, n), dtype=np.float32)
    means, fcovs = [], []
    for f in range(F):
        b, ctr = base[f % 8], traj[:, f % traj.shape[1]]
        x[f] = b[0] + ctr.astype(np.float32)[:, None]
        means.append(b[2] + ctr[None, :])
        fcovs.append(b[3])
    return x, np.stack(means), np.stack(fcovs)


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
