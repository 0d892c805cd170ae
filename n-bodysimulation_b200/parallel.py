"""parallel.py -- multi-GPU orchestration: one process per GPU, torch.distributed for
the exchange (NCCL over NVLink on GPUs; gloo in the CPU tests of this logic).

The reference is single-GPU (device 0, default stream, no
collectives).  Sharding per BASELINE.json's north_star: rank g owns N/G target
bodies; every step all-gathers the packed positions, overlapped with the
interactions against the rank's own slice:

    compute stream :  local(own x own) ------------------> remote(own x others)+integrate
    comm stream    :  all_gather(new positions, in place) -^   (event)

`exchange="p2p"` replaces the gather by a FUSED exchange: every rank maps its peers' source
buffers (CUDA IPC handles passed around with all_gather_object) and the step kernel's epilogue
stores each new position straight into every rank's next buffer over NVLink -- one kernel per
step over all sources.  The cross-rank barrier between steps is inside the kernel (flags in peer
memory; `barrier="device"`, chosen by "auto" when every rank has its own GPU) or a 4-byte all-reduce
on the same stream (`barrier="collective"`):

    compute stream :  fused step (own targets x all sources; epilogue -> own + peers' next buffers) [| all-reduce]

Body order and slot geometry come from the C ABI (nbody_shard_sources): rank r's
slice sits at byte offset r*slot_bytes of the packed source buffer, so the
in-place all-gather (`input` = own slice view of `output`) needs no staging copy.
"""
from __future__ import annotations

import numpy as np


class DevBuf:
    """a raw device allocation exposed through __cuda_array_interface__ (uint8), so
    torch can view memory owned by the C library without copying"""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


def as_tensor(ptr: int, nbytes: int, device):
    import torch
    return torch.as_tensor(DevBuf(ptr, nbytes), device=device)


def _device_identity(torch, device):
    """(host, GPU uuid): equal on two ranks exactly when they run on the same physical GPU"""
    import socket
    props = torch.cuda.get_device_properties(device)
    return (socket.gethostname(), str(getattr(props, "uuid", torch.device(device).index)))


class ShardedSimulation:
    """Steps a sharded run.  `sim` is this rank's nb.Simulation(rank=r, world=G) -- or any
    object with shard_local_async / shard_remote_async / shard_sources (the gloo CPU test
    passes a host-memory stand-in).  `group` is a torch.distributed process group."""

    def __init__(self, sim, rank: int, world: int, group=None, device=None, overlap: bool = True, exchange: str = "nccl",
                 barrier: str = "auto"):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.sim, self.rank, self.world, self.group = sim, rank, world, group
        self.device = device
        self.on_gpu = device is not None and torch.device(device).type == "cuda"
        self.overlap = overlap and self.on_gpu
        self._views = {}
        if self.on_gpu:
            # always an explicit torch stream: the C ABI maps stream 0 to the context's own non-blocking stream, which
            # nothing orders against torch's collectives.  overlap=False: ONE stream for kernels and collectives
            # (fully serialised); overlap=True: the gather runs on a second stream beside the own-slice kernel.
            self.compute = torch.cuda.Stream(device=device)
            self.comm = torch.cuda.Stream(device=device) if self.overlap else self.compute
        self.steps_done = 0
        assert exchange in ("nccl", "p2p")
        assert barrier in ("auto", "device", "collective")
        self.exchange = exchange if world > 1 else "nccl"
        self.barrier = None
        if self.exchange == "p2p":
            on_gpu = device is not None and torch.device(device).type == "cuda"
            # every rank opens every other rank's two source buffers (cudaIpcOpenMemHandle)
            mine = (sim.shard_ipc_handles(), _device_identity(torch, device) if on_gpu else None)
            got = [None] * world
            dist.all_gather_object(got, mine, group=group)
            for r in range(world):
                if r != rank:
                    sim.shard_attach_ipc(r, got[r][0])
            # kernels that wait for each other must not share a GPU
            own_gpu_each = on_gpu and len({g[1] for g in got}) == world
            self._flag = torch.zeros(1, dtype=torch.int32, device=device if device is not None else "cpu")
            # cross-rank barrier between fused steps: inside the step kernel itself (flags in peer memory, no
            # collective call per step) where the engine offers it, else a 4-byte all-reduce on the step's stream
            self.barrier = "collective"
            if barrier != "collective" and own_gpu_each and hasattr(sim, "shard_set_device_barrier"):
                try:
                    sim.shard_set_device_barrier(True)
                    ok = 1
                except Exception:  # noqa: BLE001  (small-N family kernels have no in-kernel barrier)
                    ok = 0
                agree = torch.tensor([ok], dtype=torch.int32, device=device)
                dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=group)
                if int(agree.item()) == 1:
                    self.barrier = "device"
                elif ok:
                    sim.shard_set_device_barrier(False)
            if barrier == "device" and self.barrier != "device":
                raise RuntimeError("in-kernel barrier needs every rank on its own GPU and a kernel family that has it")
            # nobody steps before everybody has uploaded and attached
            dist.all_reduce(self._flag, group=group)
            if on_gpu:
                torch.cuda.synchronize(device)

    def _source_views(self, which: int):
        ptr, total, slot = self.sim.shard_sources(which)
        key = (ptr.ctypes.data if isinstance(ptr, np.ndarray) else ptr, total)
        if key not in self._views:
            if isinstance(ptr, np.ndarray):          # CPU stand-in hands out a numpy buffer
                full = self.torch.from_numpy(ptr)
            else:
                full = as_tensor(ptr, total, self.device)
            self._views[key] = (full, full[self.rank * slot:(self.rank + 1) * slot])
        return self._views[key]

    def gather_current(self):
        """all-gather the CURRENT source buffer in place (own slice -> everyone)"""
        full, own = self._source_views(0)
        return self.dist.all_gather_into_tensor(full, own, group=self.group, async_op=False)

    def step(self) -> None:
        """one step; on return the current source buffer holds everyone's new positions
        (work may still be in flight on self.compute / self.comm when overlapping)"""
        if self.world == 1:
            self.sim.shard_local_async(self.compute.cuda_stream if self.on_gpu else 0)
            self.steps_done += 1
            return
        if self.exchange == "p2p":
            torch = self.torch
            # barrier after the step: my stores have left (kernel complete) and every peer's have landed
            # before the next step reads them; also nobody overwrites a buffer a peer is still reading
            if not self.on_gpu:
                self.sim.shard_step_fused_async(0)
                self.dist.all_reduce(self._flag, group=self.group)
            else:
                st = self.compute
                self.sim.shard_step_fused_async(st.cuda_stream)
                if self.barrier != "device":
                    with torch.cuda.stream(st):
                        self.dist.all_reduce(self._flag, group=self.group)
            self.steps_done += 1
            return
        if not self.on_gpu:
            self.sim.shard_local_async(0)
            self.sim.shard_remote_async(0)
            self.gather_current()
            self.steps_done += 1
            return
        torch = self.torch
        if not self.overlap:   # one stream: local, remote, gather in program order
            self.sim.shard_local_async(self.compute.cuda_stream)
            self.sim.shard_remote_async(self.compute.cuda_stream)
            with torch.cuda.stream(self.compute):
                self.gather_current()
            self.steps_done += 1
            return
        # 1. own x own on the compute stream; needs only this rank's slice, which the previous
        #    remote kernel (same stream) wrote -- so it runs while the gather is in flight
        self.sim.shard_local_async(self.compute.cuda_stream)
        # 2. own x others once the gather of the previous step has landed
        self.compute.wait_stream(self.comm)
        self.sim.shard_remote_async(self.compute.cuda_stream)
        # 3. publish the new positions: in-place all-gather of the (now current) buffer
        self.comm.wait_stream(self.compute)
        with torch.cuda.stream(self.comm):
            self.gather_current()
        self.steps_done += 1

    def synchronize(self) -> None:
        if self.on_gpu:
            self.compute.synchronize()
            self.comm.synchronize()

    def quiesce(self) -> None:
        """every rank has finished everything it launched and nobody has started anything new: required around
        nbody_upload / nbody_snapshot_load on a context whose buffers peers store into (fused exchange)"""
        self.synchronize()
        if self.world > 1:
            self.dist.barrier(group=self.group)
            if self.on_gpu:
                self.torch.cuda.synchronize(self.device)
