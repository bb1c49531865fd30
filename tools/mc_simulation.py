#!/usr/bin/env python3
"""Headless Python-3 twin of the reference driver (MCSimulation.py:100-270): same command line
("MC" or "GMM"), same command sequence to the module, same 200-run loop, same checkpoint_*.txt /
simReport_*.txt / GMMcheckpoint_*.txt / GMMsimReport_*.txt files (:16-77, :226-256).  What differs: no
OpenRAVE, no viewer -- the module is pcoll_b200.MCModule (libmcsim.so -> libmcsim_cuda.so), which
extracts the collision geometry from the env XML itself.

  python tools/mc_simulation.py MC  [--env data/scenes/pr2test2.env.xml] [--num-simulations 200]
  python tools/mc_simulation.py GMM [--num-particles 10000] [--outdir reports] [--mode shared|philox]
"""
import argparse
import datetime
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402


def _stamp():
    return datetime.datetime.fromtimestamp(time.time()).strftime("%Y-%m-%d_%H_%M_%S")


def write_report(path, gmm, envfile, plan, num_particles, num_simulations, sim_times, proportions, num_gaussians, num_samples):
    """simReport_*.txt (MCSimulation.py:16-44) / GMMsimReport_*.txt (:47-77): same lines, same order."""
    with open(path, "w") as f:
        f.write("Environment: %s\n" % envfile)
        f.write("Num Landmarks: %d\n" % plan["landmarks"].shape[1])
        f.write("Landmarks: \n%s\n" % plan["landmarks"])
        f.write("Alphas: \n%s\n" % pcoll_b200.plugin.list2string(plan["alphas"]))
        f.write("Sensor Noise Variance: %s\n" % plan["Q"])
        f.write("Initial Covariance: \n%s\n" % np.asarray(plan["cov0"]))
        f.write("---------------------------------\n")
        f.write("NumSimulations: %d\n" % num_simulations)
        if gmm:
            f.write("Num Samples: %d\n" % num_samples)
            f.write("Num Gaussians: %d\n" % num_gaussians)
        else:
            f.write("Num Particles: %d\n" % num_particles)
        f.write("Simulation Times: \n%s\n" % sim_times)
        f.write("Collision Proportions: \n%s\n" % proportions)
        f.write("Average Sim Time: %s\n" % np.average(sim_times))
        f.write("Average Prob Collision: %s\n" % np.average(proportions))
        f.write("---------------------------------\n")
        f.write("Trajectory: \n%s\n" % plan["trajectory"].transpose())
        f.write("Odometry: \n%s\n" % plan["odometry"].transpose())


def main(argv=None):
    ap = argparse.ArgumentParser(usage='%(prog)s "MC" or "GMM" [options]')
    ap.add_argument("simoption", choices=["MC", "GMM"])
    ap.add_argument("--env", default=os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml"))
    ap.add_argument("--trajectory", default=os.path.join(ROOT, "data", "plans", "trajectory.dat"))
    ap.add_argument("--odometry", default=os.path.join(ROOT, "data", "plans", "odometry.dat"))
    ap.add_argument("--num-simulations", type=int, default=200)     # MCSimulation.py:214
    ap.add_argument("--num-particles", type=int, default=10000)     # :164
    ap.add_argument("--num-gaussians", type=int, default=3)         # :204
    ap.add_argument("--mode", default="shared", choices=["shared", "philox"])
    ap.add_argument("--seed", type=int, default=None)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--outdir", default=".")
    ap.add_argument("--batch", action="store_true",
                    help="GMM only: the whole run loop as ONE runGMMEstimationBatch command (same runs, same files; "
                         "simTime of each run = batch time / runs)")
    a = ap.parse_args(argv)

    plan = pcoll_b200.plan_from_pickles(a.trajectory, a.odometry)
    mod = pcoll_b200.MCModule(a.env, device=a.device)
    print(mod.SendCommand("help"))
    print(mod.SendCommand("MyCommand [0.1,4.5,7.5,4.7]"))
    gmm = a.simoption == "GMM"
    pcoll_b200.push_parameters(mod, plan, num_particles=a.num_particles,
                               num_gaussians=a.num_gaussians if gmm else None,
                               num_gmm_samples=a.num_particles if gmm else None)   # :205: numGMMSamples = numParticles
    mod.SendCommand("setMode " + a.mode)
    if a.seed is not None:
        mod.SendCommand("setSeed %d" % a.seed)

    os.makedirs(a.outdir, exist_ok=True)
    st = _stamp()
    sim_times, proportions = [], []
    ck = os.path.join(a.outdir, ("GMMcheckpoint_" if gmm else "checkpoint_") + st + ".txt")
    with open(ck, "w") as f2:
        batch = None
        if gmm and a.batch:
            start = time.perf_counter()
            batch = [float(v) for v in mod.SendCommand("runGMMEstimationBatch %d" % a.num_simulations).split()]
            batch_time = (time.perf_counter() - start) / a.num_simulations
        for i in range(a.num_simulations):
            start = time.perf_counter()
            collprop = batch[i] if batch else float(mod.SendCommand("runGMMEstimation" if gmm else "runSimulation"))
            sim_time = batch_time if batch else time.perf_counter() - start
            print("Python got collision proportion: ", collprop)
            print("Python Simulation Time: ", sim_time)
            sim_times.append(sim_time)
            proportions.append(collprop)
            f2.write("Simulation: %d\nsimTime: %s\ncollProp: %s\n" % (i, sim_time, collprop))
            f2.flush()
            os.fsync(f2.fileno())
    print("SimTimes: \n", sim_times)
    print("Collision Probs: \n", proportions)
    rp = os.path.join(a.outdir, ("GMMsimReport_" if gmm else "simReport_") + _stamp() + ".txt")
    write_report(rp, gmm, a.env, plan, a.num_particles, a.num_simulations, sim_times, proportions, a.num_gaussians, a.num_particles)
    mod.close()
    return dict(checkpoint=ck, report=rp, proportions=proportions, sim_times=sim_times)


if __name__ == "__main__":
    main()
