"""CPU test (gloo, world_size 2 and 3) of the multi-GPU host logic in
n-bodysimulation_b200/parallel.py: slot geometry from the C ABI, in-place
all-gather of each rank's slice, local -> remote -> gather ordering.  The
per-rank kernels are replaced by a host-memory stand-in that computes with the
oracle (tests only); the sharded result must equal the oracle's unsharded step
bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class HostStandIn:
    """same interface as nb.Simulation's shard_* methods, state in numpy, math by the oracle"""

    def __init__(self, nb, oracle, x0, rank, world, eps2=0.002, dt=0.1):
        self.o, self.rank, self.world, self.eps2, self.dt = oracle, rank, world, eps2, dt
        n = len(x0)
        self.n = n
        self.per = ((n + world - 1) // world + 1) // 2 * 2
        self.slot = nb.lib.nbody_shard_slot_bytes(n, world)
        self.begin = rank * self.per
        self.count = max(0, min(n - self.begin, self.per))
        self.buf = [np.zeros(self.slot * world, np.uint8), np.zeros(self.slot * world, np.uint8)]
        self.cur = 0
        for r in range(world):                                   # like nbody_upload: every rank's slice in its slot
            b = r * self.per
            c = max(0, min(n - b, self.per))
            self._slot_view(0, r)[:c] = x0[b:b + c]
        self.V = np.zeros((self.count, 4), np.float32)
        self.A = np.zeros((self.count, 4), np.float32)
        self.calls = []

    def _slot_view(self, which, r):
        raw = self.buf[self.cur ^ which][r * self.slot:(r + 1) * self.slot]
        return raw.view(np.float32).reshape(-1, 4)

    def _all_bodies(self):
        parts = []
        for r in range(self.world):
            c = max(0, min(self.n - r * self.per, self.per))
            parts.append(self._slot_view(0, r)[:c])
        return np.ascontiguousarray(np.concatenate(parts))

    def shard_sources(self, which=0):
        return self.buf[self.cur ^ which], self.slot * self.world, self.slot

    def shard_local_async(self, stream=0):
        self.calls.append("local")

    def shard_remote_async(self, stream=0):
        self.calls.append("remote")
        X = self._all_bodies()
        idx = np.arange(self.begin, self.begin + self.count, dtype=np.int32)
        acc = self.o.accel_f32(X, idx, self.eps2)
        self.A[:, :3] = acc
        hdt = np.float32(0.5) * np.float32(self.dt)
        own = X[idx].copy()
        self.V[:, :3] = self.V[:, :3] + hdt * acc
        own[:, :3] = own[:, :3] + np.float32(self.dt) * self.V[:, :3]
        self._slot_view(1, self.rank)[:self.count] = own
        self.cur ^= 1


def _worker(rank, world, port, n, steps, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch.distributed as dist
    from importlib import import_module
    import nbody_b200 as nb
    import oracle
    par = import_module("n-bodysimulation_b200.parallel")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    x0 = nb.ic_uniform(n, seed=99)
    sim = HostStandIn(nb, oracle, x0, rank, world)
    runner = par.ShardedSimulation(sim, rank, world, device=None)
    for _ in range(steps):
        runner.step()
    runner.quiesce()      # the fence bench.py puts around re-uploads: synchronise + cross-rank barrier (no streams on CPU)
    assert sim.calls == ["local", "remote"] * steps
    np.save(os.path.join(out_dir, f"x_{rank}.npy"), sim._all_bodies())       # everyone must hold all new positions
    np.save(os.path.join(out_dir, f"v_{rank}.npy"), sim.V)
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 1001), (3, 700)])
def test_sharded_runner_equals_unsharded_oracle(tmp_path, oracle, nb, world, n):
    steps = 2
    port = 29500 + os.getpid() % 500 + world
    mp.spawn(_worker, args=(world, port, n, steps, str(tmp_path)), nprocs=world, join=True)
    x0 = nb.ic_uniform(n, seed=99)
    X, V, A = x0.copy(), np.zeros_like(x0), np.zeros_like(x0)
    for _ in range(steps):
        oracle.step_snapshot(X, A, V)
    per = ((n + world - 1) // world + 1) // 2 * 2
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"x_{r}.npy"), X), r        # bit exact, on every rank
        assert np.array_equal(np.load(tmp_path / f"v_{r}.npy")[:, :3], V[r * per:(r + 1) * per, :3]), r


# ---- exchange="p2p": peers' buffers mapped, the step stores into all of them, tiny all-reduce as barrier ----
class SharedMemStandIn(HostStandIn):
    """HostStandIn whose two source buffers live in POSIX shared memory, so that other ranks
    (processes) can map them: the CPU twin of the CUDA IPC handles + P2P stores of the fused exchange.
    The 128-byte 'handle' carries the two segment names."""

    def __init__(self, nb, oracle, x0, rank, world, **kw):
        from multiprocessing import shared_memory
        super().__init__(nb, oracle, x0, rank, world, **kw)
        self._shm = [shared_memory.SharedMemory(create=True, size=self.slot * world) for _ in range(2)]
        for i in range(2):
            mapped = np.ndarray(self.slot * world, np.uint8, buffer=self._shm[i].buf)
            mapped[:] = self.buf[i]
            self.buf[i] = mapped
        self.peer, self._peer_shm = {}, []

    def shard_ipc_handles(self):
        return ("|".join(s.name for s in self._shm)).encode().ljust(128, b"\0")

    def shard_attach_ipc(self, peer, handles):
        from multiprocessing import shared_memory
        assert peer != self.rank and peer not in self.peer and len(handles) == 128
        segs = [shared_memory.SharedMemory(name=nm) for nm in handles.rstrip(b"\0").decode().split("|")]
        self._peer_shm += segs
        self.peer[peer] = [np.ndarray(self.slot * self.world, np.uint8, buffer=s.buf) for s in segs]

    def shard_step_fused_async(self, stream=0):
        assert len(self.peer) == self.world - 1, "all peers must be attached"
        self.calls.append("fused")
        nxt = self.cur ^ 1
        self.shard_remote_async()                                # computes, writes own next slot, flips
        self.calls.pop()
        own = self.buf[nxt][self.rank * self.slot:(self.rank + 1) * self.slot]
        for bufs in self.peer.values():                          # the "P2P stores"
            bufs[nxt][self.rank * self.slot:(self.rank + 1) * self.slot] = own

    def close(self):
        self.buf = [None, None]
        self.peer = {}
        for s in self._peer_shm:
            s.close()
        for s in self._shm:
            s.close()
            s.unlink()


def _worker_p2p(rank, world, port, n, steps, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch.distributed as dist
    from importlib import import_module
    import nbody_b200 as nb
    import oracle
    par = import_module("n-bodysimulation_b200.parallel")
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    sim = SharedMemStandIn(nb, oracle, nb.ic_uniform(n, seed=99), rank, world)
    runner = par.ShardedSimulation(sim, rank, world, device=None, exchange="p2p")
    assert runner.barrier == "collective"                        # no GPUs: "auto" must not pick the in-kernel barrier
    try:
        par.ShardedSimulation(SharedMemStandIn.__new__(SharedMemStandIn), rank, world, device=None, exchange="nccl", barrier="bogus")
        raise SystemExit("bad barrier name accepted")
    except AssertionError:
        pass
    for _ in range(steps):
        runner.step()                                            # ends with the all-reduce barrier
    assert sim.calls == ["fused"] * steps
    np.save(os.path.join(out_dir, f"x_{rank}.npy"), sim._all_bodies())
    np.save(os.path.join(out_dir, f"v_{rank}.npy"), sim.V)
    dist.barrier()                                               # nobody unmaps while a peer may still read
    sim.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 1001), (3, 700)])
def test_fused_exchange_runner_equals_unsharded_oracle(tmp_path, oracle, nb, world, n):
    steps = 3
    port = 29600 + os.getpid() % 300 + world
    mp.spawn(_worker_p2p, args=(world, port, n, steps, str(tmp_path)), nprocs=world, join=True)
    x0 = nb.ic_uniform(n, seed=99)
    X, V, A = x0.copy(), np.zeros_like(x0), np.zeros_like(x0)
    for _ in range(steps):
        oracle.step_snapshot(X, A, V)
    per = ((n + world - 1) // world + 1) // 2 * 2
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"x_{r}.npy"), X), r
        assert np.array_equal(np.load(tmp_path / f"v_{r}.npy")[:, :3], V[r * per:(r + 1) * per, :3]), r
