"""e2e_breakdown.py -- where the host-buffer step spends its time (run on the GPU box)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import nbody_b200 as nb
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
x0 = nb.ic_uniform(n, seed=1)
hx = torch.from_numpy(x0).pin_memory(); hv = torch.zeros(n, 4).pin_memory()
ox = torch.empty(n, 4).pin_memory(); ov = torch.empty(n, 4).pin_memory(); oa = torch.empty(n, 4).pin_memory()
sim = nb.Simulation(n)
for it in range(4):
    t0 = time.perf_counter(); nb._check(nb.lib.nbody_upload(sim._h, hx.data_ptr(), hv.data_ptr()))
    t1 = time.perf_counter(); sim.step(1)
    t2 = time.perf_counter(); nb._check(nb.lib.nbody_download(sim._h, ox.data_ptr(), ov.data_ptr(), oa.data_ptr()))
    t3 = time.perf_counter()
    print(f"iter {it}: upload {1e3*(t1-t0):.2f} ms  step {1e3*(t2-t1):.2f} ms (device {sim.elapsed_ms():.2f})  download {1e3*(t3-t2):.2f} ms")
