"""n-bodysimulation_b200 -- Python face of libnbody_b200.so (ctypes over the C ABI).

The product is the CUDA library; this module only loads it and mirrors the
reference's launch interface so tests and bench.py read like the reference's
callers (TestProject/main.cpp:142-160, TestProject/validation.cpp:55-103):

    simulate(d_bodies, d_accelerations, d_velocity, N)   # kernel.cuh:2
    Simulation(n, ...).upload(x, v).step(k).download()   # resident state

There is NO CPU fallback: if the shared library is missing the import fails,
and every compute call fails without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# NBODY_LIB=libnbody_b200_sweep.so selects the sweep build (`make -C n-bodysimulation_b200 sweep`: every rejected
# tile shape of round 1 as an extra variant) for tools/sweep.py; the default is the product library
LIB_PATH = os.path.join(_HERE, os.environ.get("NBODY_LIB", "libnbody_b200.so"))

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
        "(there is no CPU fallback for the N-body step)")

lib = C.CDLL(LIB_PATH)

NBODY_INTEG_F32, NBODY_INTEG_F64, NBODY_INTEG_MIXED = 0, 1, 2
NBODY_EXCHANGE_P2P, NBODY_EXCHANGE_NCCL, NBODY_EXCHANGE_P2P_EVENTS = 0, 1, 2
NBODY_MAX_DEVICES = 16
NBODY_SUM_TWO_LEVEL, NBODY_SUM_SINGLE = 0, 1


class NBodyParams(C.Structure):
    """struct nbody_params of include/nbody_b200.h (TestProject/constants.h at run time)."""
    _fields_ = [("n_bodies", C.c_int), ("rank", C.c_int), ("world", C.c_int),
                ("eps2", C.c_float), ("dt", C.c_float), ("integrator", C.c_int),
                ("device", C.c_int), ("variant", C.c_int), ("grid_ctas", C.c_int),
                ("eps2_f64", C.c_double), ("dt_f64", C.c_double),
                ("deterministic", C.c_int), ("n_devices", C.c_int), ("exchange", C.c_int),
                ("devices", C.c_int * NBODY_MAX_DEVICES), ("summation", C.c_int)]


class NBodyStepOpts(C.Structure):
    """struct nbody_step_opts: the reference's compile-time switches (constants.h:4-7) at run time."""
    _fields_ = [("eps2", C.c_float), ("dt", C.c_float), ("integrator", C.c_int), ("va_float3", C.c_int), ("variant", C.c_int)]


class NBodyView(C.Structure):
    """struct nbody_view: window size and shader uniforms of simulation_visualization.cpp:203-207."""
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("max_x", C.c_float), ("max_y", C.c_float), ("max_z", C.c_float),
                ("min_w", C.c_float), ("max_w", C.c_float), ("depth_test", C.c_int)]


_fp = C.POINTER(C.c_float)
_vp = C.c_void_p
# every symbol include/nbody_b200.h declares: (restype, argtypes)
SIGNATURES = {
    "nbody_params_default": (C.c_int, [C.POINTER(NBodyParams)]),
    "nbody_create": (C.c_int, [C.POINTER(_vp), C.POINTER(NBodyParams)]),
    "nbody_destroy": (None, [_vp]),
    "nbody_upload": (C.c_int, [_vp, _vp, _vp]),
    "nbody_owned_range": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "nbody_download": (C.c_int, [_vp, _vp, _vp, _vp]),
    "nbody_step": (C.c_int, [_vp, C.c_int]),
    "nbody_step_async": (C.c_int, [_vp, C.c_int]),
    "nbody_sync": (C.c_int, [_vp]),
    "nbody_elapsed_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "nbody_kernel_clock_mhz": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "nbody_energy": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "nbody_shard_local_async": (C.c_int, [_vp, _vp]),
    "nbody_shard_remote_async": (C.c_int, [_vp, _vp]),
    "nbody_shard_sources": (C.c_int, [_vp, C.c_int, C.POINTER(_vp), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    "nbody_shard_slot_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "nbody_shard_ipc_handles": (C.c_int, [_vp, _vp]),
    "nbody_shard_attach_ipc": (C.c_int, [_vp, C.c_int, _vp]),
    "nbody_shard_buffers": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp)]),
    "nbody_shard_attach_ptr": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "nbody_shard_step_fused_async": (C.c_int, [_vp, _vp]),
    "nbody_shard_set_device_barrier": (C.c_int, [_vp, C.c_int]),
    "nbody_shard_barrier_timeouts": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "nbody_shard_barrier_wait_ns": (C.c_int, [_vp, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong), C.c_int]),
    "nbody_set_devices": (C.c_int, [C.POINTER(C.c_int), C.c_int]),
    "nbody_group_info": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "nbody_step_inplace_f4": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_float, C.c_float]),
    "nbody_step_inplace_noalloc_f4": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_float, C.c_float, _vp]),
    "nbody_noalloc_max_bodies": (C.c_int, []),
    "nbody_step_host_f3": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_double, C.c_double]),
    "nbody_step_opts_default": (C.c_int, [C.POINTER(NBodyStepOpts)]),
    "nbody_step_inplace_opts": (C.c_int, [_vp, _vp, _vp, C.c_int, C.POINTER(NBodyStepOpts)]),
    "nbody_release_cached": (None, []),
    "nbody_view_default": (C.c_int, [C.POINTER(NBodyView)]),
    "nbody_render_device": (C.c_int, [_vp, C.c_int, C.POINTER(NBodyView), _vp]),
    "nbody_render": (C.c_int, [_vp, C.POINTER(NBodyView), _vp]),
    "nbody_write_ppm": (C.c_int, [C.c_char_p, _vp, C.c_int, C.c_int]),
    "nbody_snapshot_save": (C.c_int, [_vp, C.c_char_p]),
    "nbody_snapshot_load": (C.c_int, [_vp, C.c_char_p]),
    "nbody_snapshot_info": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_longlong), C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "nbody_steps_done": (C.c_int, [_vp, C.POINTER(C.c_longlong)]),
    "nbody_last_error": (C.c_char_p, []),
    "nbody_launch_count": (C.c_longlong, []),
    "nbody_variant_info": (C.c_int, [C.c_int, C.c_char_p, C.c_size_t]),
    "nbody_variant_id": (C.c_int, [C.c_int]),
    "nbody_choose_variant": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "nbody_autotune_report": (C.c_int, [C.c_char_p, C.c_size_t]),
    "nbody_ctx_info": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "nbody_ctx_fixup": (C.c_int, [_vp, C.POINTER(C.c_int)]),
    "nbody_ic_uniform": (C.c_int, [_vp, C.c_int, C.c_ulonglong, C.c_float, C.c_float, C.c_float]),
    "nbody_ic_plummer": (C.c_int, [_vp, _vp, C.c_int, C.c_ulonglong]),
}
for _name, (_res, _args) in SIGNATURES.items():
    _f = getattr(lib, _name)
    _f.restype, _f.argtypes = _res, _args


class NBodyError(RuntimeError):
    """Python twin of the std::runtime_error the reference's launcher throws (kernel.cu:634,640)."""


def _check(rc: int) -> None:
    if rc != 0:
        raise NBodyError(f"[{rc}] {lib.nbody_last_error().decode()}")


def _ptr(a) -> int:
    """address of a C-contiguous float32 numpy array, a torch tensor, or a raw int device pointer"""
    if a is None:
        return 0
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        assert a.dtype in (np.float32, np.uint8) and a.flags["C_CONTIGUOUS"], "need a contiguous float32 (or uint8 frame) array"
        return a.ctypes.data
    return a.data_ptr()   # torch tensor


def launch_count() -> int:
    return int(lib.nbody_launch_count())


def variants() -> dict:
    buf = C.create_string_buffer(320)
    n = lib.nbody_variant_info(1, buf, 320)
    out = {}
    for k in range(n):
        i = lib.nbody_variant_id(k)
        lib.nbody_variant_info(i, buf, 320)
        out[i] = buf.value.decode()
    return out


def choose_variant(n_targets: int, n_sources: int, sm_count: int = 148):
    """(variant id, measured) NBODY_VARIANT_AUTO resolves to by the static rules; measured = True: one of the
    equivalent large-N kernels, timed on the device when the first such context is created.  Host-only."""
    m = C.c_int(0)
    vid = lib.nbody_choose_variant(n_targets, n_sources, sm_count, C.byref(m))
    if vid == 0:
        raise NBodyError(f"[-1] {lib.nbody_last_error().decode()}")   # NBODY_E_INVALID
    return vid, bool(m.value)


def autotune_report() -> str:
    buf = C.create_string_buffer(512)
    _check(lib.nbody_autotune_report(buf, 512))
    return buf.value.decode()


# ---- the reference's launch interface ------------------------------------------------
_eps2, _dt = 0.002, 0.1   # TestProject/constants.h:25-26


def simulate_set_parameters(eps2: float, dt: float) -> None:
    global _eps2, _dt
    _eps2, _dt = float(eps2), float(dt)


def set_devices(devices=None) -> None:
    """nbody_set_devices: spread every following simulate() / simulate_opts() call over these GPUs (first = the
    device holding the arrays); None or one device: single-device behaviour (the reference's, kernel.cu:630)"""
    if not devices or len(devices) <= 1:
        _check(lib.nbody_set_devices(None, 0))
    else:
        arr = (C.c_int * len(devices))(*devices)
        _check(lib.nbody_set_devices(arr, len(devices)))


def simulate(d_bodies, d_accelerations, d_velocity, N: int) -> None:
    """TestProject/kernel.cuh:2 -- DEVICE float4 arrays (torch CUDA tensors or raw
    pointers), one step in place, synchronous."""
    _check(lib.nbody_step_inplace_f4(_ptr(d_bodies), _ptr(d_accelerations), _ptr(d_velocity), int(N), _eps2, _dt))


def simulate_noalloc(d_bodies, d_accelerations, d_velocity, N: int, stream=None) -> None:
    """nbody_step_inplace_noalloc_f4: the same step as simulate() for N <= nbody_noalloc_max_bodies(), enqueued on
    `stream` (a cudaStream_t as int, None = the legacy default stream) without any allocation: the kernel's
    scratch is a static device array of the library.  Not synchronised; overlapping calls race (see the header).
    No counterpart in the reference."""
    _check(lib.nbody_step_inplace_noalloc_f4(_ptr(d_bodies), _ptr(d_accelerations), _ptr(d_velocity), int(N), _eps2, _dt,
                                             C.c_void_p(stream) if stream else None))


def simulate_host_f3(bodies: np.ndarray, accelerations: np.ndarray, velocity: np.ndarray, N: int,
                     eps2: float = 0.002, dt: float = 0.01) -> None:
    """Sim-Without-OpenGL-Integration/kernel.cuh:5 -- HOST float4 bodies, float3 velocity."""
    _check(lib.nbody_step_host_f3(_ptr(bodies), _ptr(accelerations), _ptr(velocity), int(N), eps2, dt))


def simulate_opts(d_bodies, d_accelerations, d_velocity, N: int, *, integrator: int = NBODY_INTEG_F32,
                  va_float3: bool = False, variant: int = 0) -> None:
    """nbody_step_inplace_opts: the in-place step with the reference's build switches as options"""
    o = NBodyStepOpts()
    _check(lib.nbody_step_opts_default(C.byref(o)))
    o.eps2, o.dt, o.integrator, o.va_float3, o.variant = _eps2, _dt, integrator, int(va_float3), variant
    _check(lib.nbody_step_inplace_opts(_ptr(d_bodies), _ptr(d_accelerations), _ptr(d_velocity), int(N), C.byref(o)))


def simulateHeadless_embParallel(d_bodies, d_accelerations, d_velocity, d_reduceMatrix, N: int) -> None:
    """headless form of TestProject/kernel.cuh:4 (also _fadl / _fadl4, :6-7): float4 V/A, reduction-flavour integrator"""
    simulate_opts(d_bodies, d_accelerations, d_velocity, N, integrator=NBODY_INTEG_MIXED)


def simulateHeadless_embParallel_float3(d_bodies, d_accelerations, d_velocity, d_reduceMatrix, N: int) -> None:
    """headless form of TestProject/kernel.cuh:5: float3 V/A device arrays"""
    simulate_opts(d_bodies, d_accelerations, d_velocity, N, integrator=NBODY_INTEG_MIXED, va_float3=True)


# ---- headless frame (simulation_visualization.cpp's shaders as CUDA) ------------------------------
def default_view(**kw) -> NBodyView:
    v = NBodyView()
    _check(lib.nbody_view_default(C.byref(v)))
    for k, val in kw.items():
        setattr(v, k, val)
    return v


def render_device(d_bodies, N: int, d_rgba, view: NBodyView | None = None) -> None:
    """DEVICE float4 positions -> DEVICE uint8 RGBA frame (height, width, 4), rows top-down"""
    view = view or default_view()
    _check(lib.nbody_render_device(_ptr(d_bodies), int(N), C.byref(view), _ptr(d_rgba)))


def write_ppm(path: str, rgba: np.ndarray) -> None:
    h, w, _ = rgba.shape
    _check(lib.nbody_write_ppm(path.encode(), rgba.ctypes.data, w, h))


def snapshot_info(path: str) -> dict:
    n, st, e, d = C.c_int(), C.c_longlong(), C.c_float(), C.c_float()
    _check(lib.nbody_snapshot_info(path.encode(), C.byref(n), C.byref(st), C.byref(e), C.byref(d)))
    return {"n_bodies": n.value, "steps_done": st.value, "eps2": e.value, "dt": d.value}


# ---- initial conditions -----------------------------------------------------------------
def ic_uniform(n: int, seed: int = 12345, max_xyz: float = 1e5, min_w: float = 1e5, max_w: float = 1e9) -> np.ndarray:
    x = np.zeros((n, 4), np.float32)
    _check(lib.nbody_ic_uniform(_ptr(x), n, seed, max_xyz, min_w, max_w))
    return x


def ic_plummer(n: int, seed: int = 12345):
    x = np.zeros((n, 4), np.float32)
    v = np.zeros((n, 4), np.float32)
    _check(lib.nbody_ic_plummer(_ptr(x), _ptr(v), n, seed))
    return x, v


# ---- resident-state context ----------------------------------------------------------------
class Simulation:
    """State resident in HBM (nbody_create .. nbody_destroy)."""

    def __init__(self, n: int, eps2: float = 0.002, dt: float = 0.1, *, rank: int = 0, world: int = 1,
                 device: int = 0, variant: int = 0, grid_ctas: int = 0, integrator: int = NBODY_INTEG_F32,
                 eps2_f64: float = 0.002, dt_f64: float = 0.01, devices=None, exchange: int = NBODY_EXCHANGE_P2P,
                 deterministic: bool = False, summation: int = NBODY_SUM_TWO_LEVEL):
        """devices=[d0, d1, ...]: ONE context spanning those GPUs (nbody_params.n_devices): the same upload / step /
        download / energy / save calls then drive the whole box from this thread."""
        p = NBodyParams()
        _check(lib.nbody_params_default(C.byref(p)))
        p.n_bodies, p.rank, p.world, p.eps2, p.dt = n, rank, world, eps2, dt
        p.device, p.variant, p.grid_ctas, p.integrator = device, variant, grid_ctas, integrator
        p.eps2_f64, p.dt_f64 = eps2_f64, dt_f64
        p.deterministic, p.exchange, p.summation = int(deterministic), exchange, summation
        if devices is not None and len(devices) > 1:
            assert len(devices) <= NBODY_MAX_DEVICES
            p.n_devices = len(devices)
            for i, d in enumerate(devices):
                p.devices[i] = d
        self._h = _vp()
        _check(lib.nbody_create(C.byref(self._h), C.byref(p)))
        self.n = n
        b, c = C.c_int(), C.c_int()
        _check(lib.nbody_owned_range(self._h, C.byref(b), C.byref(c)))
        self.begin, self.count = b.value, c.value

    def close(self) -> None:
        if getattr(self, "_h", None) and lib is not None:
            lib.nbody_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def info(self) -> dict:
        v, g, o, s = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _check(lib.nbody_ctx_info(self._h, C.byref(v), C.byref(g), C.byref(o), C.byref(s)))
        f = C.c_int()
        _check(lib.nbody_ctx_fixup(self._h, C.byref(f)))
        return {"variant": v.value, "variant_name": variants()[v.value].split(":")[0],
                "grid_ctas": g.value, "ctas_per_sm": o.value, "sm_count": s.value, "spread_fixup": bool(f.value)}

    def upload(self, x4: np.ndarray, v4: np.ndarray | None = None) -> "Simulation":
        assert x4.shape == (self.n, 4)
        _check(lib.nbody_upload(self._h, _ptr(x4), _ptr(v4)))
        return self

    def step(self, nsteps: int = 1) -> "Simulation":
        _check(lib.nbody_step(self._h, nsteps))
        return self

    def step_async(self, nsteps: int = 1) -> "Simulation":
        _check(lib.nbody_step_async(self._h, nsteps))
        return self

    def sync(self) -> "Simulation":
        _check(lib.nbody_sync(self._h))
        return self

    def elapsed_ms(self) -> float:
        ms = C.c_float()
        _check(lib.nbody_elapsed_ms(self._h, C.byref(ms)))
        return ms.value

    def download(self):
        x = np.empty((self.count, 4), np.float32)
        v = np.empty((self.count, 4), np.float32)
        a = np.empty((self.count, 4), np.float32)
        _check(lib.nbody_download(self._h, _ptr(x), _ptr(v), _ptr(a)))
        return x, v, a

    def kernel_clock_mhz(self) -> float:
        mhz = C.c_float()
        _check(lib.nbody_kernel_clock_mhz(self._h, C.byref(mhz)))
        return mhz.value

    def energy(self):
        ke, pe = C.c_double(), C.c_double()
        _check(lib.nbody_energy(self._h, C.byref(ke), C.byref(pe)))
        return ke.value, pe.value

    def render(self, view: NBodyView | None = None) -> np.ndarray:
        """frame of the current positions as a host uint8 array (height, width, 4)"""
        view = view or default_view()
        img = np.empty((view.height, view.width, 4), np.uint8)
        _check(lib.nbody_render(self._h, C.byref(view), img.ctypes.data))
        return img

    def save(self, path: str) -> "Simulation":
        _check(lib.nbody_snapshot_save(self._h, path.encode()))
        return self

    def load(self, path: str) -> "Simulation":
        _check(lib.nbody_snapshot_load(self._h, path.encode()))
        return self

    def steps_done(self) -> int:
        st = C.c_longlong()
        _check(lib.nbody_steps_done(self._h, C.byref(st)))
        return st.value

    # sharded stepping (one Simulation per rank; the caller owns the exchange)
    def shard_local_async(self, stream: int = 0) -> None:
        _check(lib.nbody_shard_local_async(self._h, stream))

    def shard_remote_async(self, stream: int = 0) -> None:
        _check(lib.nbody_shard_remote_async(self._h, stream))

    # fused exchange: the epilogue stores new positions into every peer's next buffer (P2P)
    def shard_ipc_handles(self) -> bytes:
        buf = C.create_string_buffer(128)
        _check(lib.nbody_shard_ipc_handles(self._h, buf))
        return buf.raw

    def shard_attach_ipc(self, peer: int, handles: bytes) -> None:
        assert len(handles) == 128
        _check(lib.nbody_shard_attach_ipc(self._h, peer, C.create_string_buffer(handles, 128)))

    def shard_buffers(self):
        a, b = _vp(), _vp()
        _check(lib.nbody_shard_buffers(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def shard_attach_ptr(self, peer: int, buf0: int, buf1: int) -> None:
        _check(lib.nbody_shard_attach_ptr(self._h, peer, buf0, buf1))

    def shard_set_device_barrier(self, enable: bool = True) -> None:
        _check(lib.nbody_shard_set_device_barrier(self._h, int(enable)))

    def shard_barrier_timeouts(self) -> int:
        t = C.c_int()
        _check(lib.nbody_shard_barrier_timeouts(self._h, C.byref(t)))
        return t.value

    def shard_barrier_wait_ns(self, reset: bool = False) -> dict:
        """time this GPU's step kernels stood waiting for their peers' flags: total / longest / number of waits"""
        t, m, w = C.c_ulonglong(), C.c_ulonglong(), C.c_ulonglong()
        _check(lib.nbody_shard_barrier_wait_ns(self._h, C.byref(t), C.byref(m), C.byref(w), int(reset)))
        return {"total_ns": t.value, "max_ns": m.value, "waits": w.value}

    def group_info(self) -> dict:
        n, e, b = C.c_int(), C.c_int(), C.c_int()
        _check(lib.nbody_group_info(self._h, C.byref(n), C.byref(e), C.byref(b)))
        return {"n_devices": n.value, "exchange": {NBODY_EXCHANGE_NCCL: "nccl", NBODY_EXCHANGE_P2P_EVENTS: "p2p-events"}.get(e.value, "p2p"), "in_kernel_barrier": bool(b.value)}

    def shard_step_fused_async(self, stream: int = 0) -> None:
        _check(lib.nbody_shard_step_fused_async(self._h, stream))

    def shard_sources(self, which: int = 0):
        p, tot, slot = _vp(), C.c_size_t(), C.c_size_t()
        _check(lib.nbody_shard_sources(self._h, which, C.byref(p), C.byref(tot), C.byref(slot)))
        return p.value, tot.value, slot.value
