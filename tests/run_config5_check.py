"""run_config5_check.py -- BASELINE config 5 (N = 8 388 608) on ONE B200, run by hand (26 s of GPU per step,
too long for the `-m gpu` suite):  python tests/run_config5_check.py  -> one JSON line.
Size-independent properties: sampled targets against the checker's fp64 sum (with the reference's own fp32
error beside it), momentum balance sum m_i a_i = 0, masses untouched, A.w = 0."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import nbody_b200 as nb  # noqa: E402
import oracle as o       # noqa: E402  (tests/ may use the checker)

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8388608
x0 = nb.ic_uniform(n, seed=12345)
with nb.Simulation(n) as sim:
    sim.upload(x0).step(1)
    ms = sim.elapsed_ms()
    x, v, a = sim.download()
    info = sim.info()
idx = np.random.default_rng(1).choice(n, 64, replace=False).astype(np.int32)
a64 = o.accel_f64(x0, idx)
err = o.norm_rel_err(a[idx], a64)
ref_err = o.norm_rel_err(o.accel_f32(x0, idx), a64)
f = a[:, :3].astype(np.float64) * x0[:, 3:4].astype(np.float64)
res = {"n": n, "variant": info["variant_name"], "ms_per_step": round(ms, 2), "g_inter_per_s": round(n * float(n) / ms / 1e6, 1),
       "sampled_targets": len(idx), "err_vs_fp64": err, "reference_fp32_err_vs_fp64": ref_err,
       "momentum_residual_rel": float(np.linalg.norm(f.sum(0)) / np.abs(f).sum()),
       "mass_untouched": bool(np.array_equal(x[:, 3], x0[:, 3])), "a_w_zero": bool(np.all(a[:, 3] == 0))}
res["ok"] = bool(err <= max(1e-5, 2 * ref_err) and res["momentum_residual_rel"] <= 1e-5 and res["mass_untouched"] and res["a_w_zero"])
print(json.dumps(res))
sys.exit(0 if res["ok"] else 1)
