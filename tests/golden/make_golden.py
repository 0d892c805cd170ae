"""make_golden.py -- generates tests/golden/*.npz from the REFERENCE'S OWN object
code (oracle/_ref/libref_cpu.so = /root/reference/TestProject/validation.cpp and
utils.cpp compiled unchanged by oracle/Makefile).  Run in the authoring
container, where /root/reference exists:

    make -C oracle && python tests/golden/make_golden.py

The vectors pin oracle/nbody_oracle.c on machines where the reference is absent
(the GPU box).  Constants are the reference's own (EPS2 0.002f, DT 0.1f).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle as o  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
assert o.ref_lib() is not None, "oracle/_ref/libref_cpu.so missing: run `make -C oracle` where /root/reference exists"

# 1. pair interactions through bodyInteractions_CPU (validation.cpp:9-24)
rng = np.random.default_rng(20261104)
K = 256
bi = np.concatenate([rng.uniform(-1e5, 1e5, (K, 3)), rng.uniform(1e5, 1e9, (K, 1))], 1).astype(np.float32)
bj = np.concatenate([rng.uniform(-1e5, 1e5, (K, 3)), rng.uniform(1e5, 1e9, (K, 1))], 1).astype(np.float32)
bj[:32, :3] = bi[:32, :3] + rng.uniform(-0.5, 0.5, (32, 3)).astype(np.float32)   # close encounters
bj[32] = bi[32]                                                                   # coincident bodies
ai = rng.uniform(-1, 1, (K, 4)).astype(np.float32)
ai[:, 3] = 0
out = np.stack([o.ref_pair(bi[k], bj[k], ai[k]) for k in range(K)])
np.savez_compressed(os.path.join(OUT, "pairs.npz"), bi=bi, bj=bj, ai=ai, out=out)

# 2. fill_with_random4 (utils.cpp:30-37), glibc rand() default seed
np.savez_compressed(os.path.join(OUT, "ic_reference.npz"), x1024=o.ref_fill_with_random4(1024))

# 3. CPU_compute as the reference runs it (literal, in place, serial) and the
#    snapshot reading built from the same object code; BASELINE config 1 (N=1024, 10 steps)
for n, steps in ((256, 10), (1024, 10)):
    X0 = o.ref_fill_with_random4(n)
    res = {"x0": X0}
    for name, fn in (("literal", o.ref_cpu_compute), ("snapshot", o.ref_snapshot_step)):
        X, V, A = X0.copy(), np.zeros_like(X0), np.zeros_like(X0)
        fn(X, A, V)
        res[f"{name}_a1"] = A.copy()          # accelerations of the first step
        for _ in range(steps - 1):
            fn(X, A, V)
        res[f"{name}_x"], res[f"{name}_v"], res[f"{name}_a"] = X, V, A
    np.savez_compressed(os.path.join(OUT, f"steps_n{n}.npz"), steps=np.int32(steps), **res)
print("golden vectors written to", OUT)
