"""sanitize_target.py -- small run of every kernel family for compute-sanitizer (one tool per gpurun call):
   compute-sanitizer --tool memcheck python tools/sanitize_target.py
(compute-sanitizer answered "closed" on this GPU pool in round 1; run plain, it is a 2-second pass over every family.)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_b200 as nb  # noqa: E402

n = 1537                                   # ragged: not a multiple of any tile or unit
x0 = nb.ic_uniform(n, seed=1)
ids = [int(a) for a in sys.argv[1:]] or [12, 3, 90, 97, 105, 9, 93, 111, 113, 60, 64, 82, 85]
for vid in ids:                            # stream-K (64-body units: several fold groups), scalar, orderings, cluster multicast / DSMEM, small family
    with nb.Simulation(n, variant=vid) as sim:
        sim.upload(x0).step(2)
        x, v, a = sim.download()
        assert np.isfinite(a).all(), vid
        if vid == ids[0]:
            frame = sim.render()
            assert frame.shape == (600, 800, 4)
    if vid in (113, 105):                  # the single-level twins (nbody_params.summation)
        with nb.Simulation(n, variant=vid, summation=nb.NBODY_SUM_SINGLE) as sim:
            sim.upload(x0).step(2)
            assert np.isfinite(sim.download()[2]).all(), vid
dx = torch.from_numpy(x0).cuda()
dv3, da3 = torch.zeros(n, 3, device="cuda"), torch.zeros(n, 3, device="cuda")
nb.simulateHeadless_embParallel_float3(dx, da3, dv3, None, n)        # 12-byte-strided epilogue
sims = [nb.Simulation(n, rank=r, world=3) for r in range(3)]          # sharded: local/remote and fused exchange
for s in sims:
    s.upload(x0)
for r, s in enumerate(sims):
    for q, peer in enumerate(sims):
        if q != r:
            s.shard_attach_ptr(q, *peer.shard_buffers())
for s in sims:
    s.shard_step_fused_async()
torch.cuda.synchronize()
for s in sims:
    s.shard_local_async()
    s.shard_remote_async()
torch.cuda.synchronize()
for s in sims:
    s.close()
with nb.Simulation(n, devices=[0, 0, 0], exchange=nb.NBODY_EXCHANGE_P2P_EVENTS, variant=113) as grp:   # ONE context over "three devices"
    grp.upload(x0).step(3)
    assert np.isfinite(grp.download()[2]).all() and np.isfinite(grp.energy()[1])
with nb.Simulation(700) as sim:                                       # graph-replayed loop
    sim.upload(nb.ic_uniform(700, seed=2)).step(40)
print("sanitize target done")
