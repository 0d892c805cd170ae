"""sweep.py -- device-timed throughput of every kernel variant (run on the GPU box).
usage: python tools/sweep.py [N] [steps] [variant ids...]   -> one JSON line per variant"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
ids = [int(a) for a in sys.argv[3:]] or list(nb.variants())
grids = [0]
x0 = nb.ic_uniform(n, seed=12345)
peak = 148 * 128 * 2 * 1.965e9
for vid in ids:
    for g in grids:
        try:
            with nb.Simulation(n, variant=vid, grid_ctas=g) as sim:
                sim.upload(x0).step(1)
                best = 1e30
                for _ in range(2):
                    sim.step(steps)
                    best = min(best, sim.elapsed_ms() / steps)
                info = sim.info()
            gi = n * float(n) / (best * 1e-3) / 1e9
            print(json.dumps({"variant": vid, "name": info["variant_name"], "n": n, "grid": info["grid_ctas"], "ctas_per_sm": info["ctas_per_sm"],
                              "ms_per_step": round(best, 4), "g_inter_per_s": round(gi, 1), "frac_fp32_peak_1965": round(gi * 20e9 / peak, 4)}), flush=True)
        except Exception as e:  # noqa: BLE001
            print(json.dumps({"variant": vid, "error": str(e)}), flush=True)
