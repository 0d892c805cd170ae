"""ref_kernel_bench.py -- baseline B2 of BASELINE.md: the reference's own kernel.cu rebuilt for
sm_100 (oracle/_ref/libref_kernel_sm100.so: one-token fix at kernel.cu:99, nvcc -O3, no fast-math,
shipped THREADS_PER_BLOCK=32 / TILE_WIDTH_FACTOR=1), timed with CUDA events around its own
simulate() (kernel.cu:628-645).  A reported baseline, not part of the product.  Run on the GPU box."""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402  (only for the seeded ICs)

retuned = len(sys.argv) > 1 and sys.argv[1] == "tpb256"
label = ("reference kernel.cu rebuilt for sm_100, RE-TUNED to THREADS_PER_BLOCK 256 (not as shipped)" if retuned
         else "reference kernel.cu rebuilt for sm_100 (32x1 launch as shipped)")
lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_kernel_sm100_tpb256.so" if retuned else "libref_kernel_sm100.so"))
simulate = getattr(lib, "_Z8simulateP6float4S0_S0_i")
simulate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
simulate.restype = None
for n, steps in ((16384, 20), (131072, 3), (1 << 20, 1)) + (() if retuned else ((8388608, 1),)):
    x = torch.from_numpy(nb.ic_uniform(n, seed=12345)).cuda()
    v, a = torch.zeros_like(x), torch.zeros_like(x)
    try:
        if n <= 131072:
            simulate(x.data_ptr(), a.data_ptr(), v.data_ptr(), n)   # warm-up
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            simulate(x.data_ptr(), a.data_ptr(), v.data_ptr(), n)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        print(json.dumps({"baseline": label, "n": n, "ms_per_step": round(ms, 3),
                          "g_inter_per_s": round(n * float(n) / ms / 1e6, 2)}), flush=True)
    except Exception as e:  # noqa: BLE001  (the launcher throws via C++; through ctypes that aborts -- see note)
        print(json.dumps({"n": n, "error": str(e)}), flush=True)
