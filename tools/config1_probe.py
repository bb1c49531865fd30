"""Developer probe: config 1 as the reference driver runs it -- runSimulation in the default `shared` mode, microseconds per call."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200
plan = pcoll_b200.load_plan_json(os.path.join(ROOT, "tests", "golden", "plan_committed.json"))
for N in (1000, 10000):
    mod = pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml"), device=0)
    pcoll_b200.push_parameters(mod, plan, num_particles=N)
    mod.SendCommand("setSeed 1")
    for _ in range(20): mod.SendCommand("runSimulation")
    t0 = time.perf_counter()
    Ps = [float(mod.SendCommand("runSimulation")) for _ in range(200)]
    dt = (time.perf_counter() - t0) / 200
    print("N=%d: %.1f us per call, P mean %.5f" % (N, dt * 1e6, sum(Ps) / len(Ps)))
    mod.close()
