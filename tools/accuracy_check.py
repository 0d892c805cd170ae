import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import nbody_b200 as nb, oracle as o
o.set_num_threads()
res = {"lib": os.environ.get("NBODY_LIB", "default")}
for n in (16384, 131072, 1 << 20):
    x0 = nb.ic_uniform(n, seed=12345)
    with nb.Simulation(n) as sim:
        sim.upload(x0).step(1)
        ms = sim.elapsed_ms()
        x, v, a = sim.download(); name = sim.info()["variant_name"]
    idx = np.random.default_rng(3).choice(n, 512, replace=False).astype(np.int32)
    a64 = o.accel_f64(x0, idx)
    e = np.linalg.norm(a[idx, :3] - a64, axis=1) / np.linalg.norm(a64, axis=1)
    r = np.linalg.norm(o.accel_f32(x0, idx) - a64, axis=1) / np.linalg.norm(a64, axis=1)
    res[n] = {"kernel": name, "ours_max": float(e.max()), "ours_median": float(np.median(e)), "ref_fp32_max": float(r.max()), "ref_fp32_median": float(np.median(r))}
print(json.dumps(res))
