"""one.py -- run one variant for a few steps (profiling target). usage: one.py N steps variant [grid]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402

n, steps, vid = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
grid = int(sys.argv[4]) if len(sys.argv) > 4 else 0
with nb.Simulation(n, variant=vid, grid_ctas=grid) as sim:
    sim.upload(nb.ic_uniform(n, seed=12345)).step(steps)
    ms = sim.elapsed_ms() / steps
    print(sim.info(), f"{ms:.3f} ms/step", f"{n * float(n) / ms / 1e6:.1f} G inter/s")
