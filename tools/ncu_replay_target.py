"""ncu target: config 1 as the reference driver runs it -- runSimulation in the plugin's default `shared` mode (replay
kernel, one real-robot thread per run) at N = 1,000 particles, a few calls."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pcoll_b200  # noqa: E402

plan = pcoll_b200.load_plan_json(os.path.join(ROOT, "tests", "golden", "plan_committed.json"))
mod = pcoll_b200.MCModule(os.path.join(ROOT, "data", "scenes", "pr2test2.env.xml"), device=0)
pcoll_b200.push_parameters(mod, plan, num_particles=int(sys.argv[1]) if len(sys.argv) > 1 else 1000)
mod.SendCommand("setSeed 1")
for _ in range(4):
    print(mod.SendCommand("runSimulation"))
mod.close()
