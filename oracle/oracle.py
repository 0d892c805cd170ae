"""oracle.py -- TEST INFRASTRUCTURE: ctypes doors into oracle/liboracle.so (our C
restatement, oracle/nbody_oracle.c) and, when present, oracle/_ref/libref_cpu*.so
(the reference's own validation.cpp / utils.cpp object code).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product never does."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_F = C.POINTER(C.c_float)
_D = C.POINTER(C.c_double)
_I = C.POINTER(C.c_int)


def _fp(a):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_F)


def build() -> str:
    """compile liboracle.so (and _ref when /root/reference is here); returns the .so path"""
    subprocess.run(["make", "-s", "-C", HERE], check=True, capture_output=True)
    return os.path.join(HERE, "liboracle.so")


def _load():
    path = os.path.join(HERE, "liboracle.so")
    if not os.path.exists(path):
        build()
    return C.CDLL(path)


_lib = _load()
_lib.oracle_num_threads.restype = C.c_int


def num_threads() -> int:
    return int(_lib.oracle_num_threads())


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def set_num_threads(n: int | None = None) -> int:
    """OpenMP threads of every checker / reference-CPU call that follows (default: all host cores this process
    may run on, whatever OMP_NUM_THREADS says -- torchrun sets it to 1 in its workers)"""
    _lib.oracle_set_num_threads(C.c_int(n or host_cores()))
    return num_threads()


def pair(bi, bj, ai, eps2=0.002):
    ai = np.array(ai, np.float32)
    _lib.oracle_pair_f32(_fp(np.asarray(bi, np.float32)), _fp(np.asarray(bj, np.float32)), _fp(ai), C.c_float(eps2))
    return ai


def step_literal(X, A, V, eps2=0.002, dt=0.1):
    """O1: in-place CPU_compute, serial (validation.cpp:28-52). Arrays modified in place."""
    _lib.oracle_step_literal_f32(_fp(X), _fp(A), _fp(V), len(X), C.c_float(eps2), C.c_float(dt))


def step_snapshot(X, A, V, eps2=0.002, dt=0.1):
    """O2 (parity target): same arithmetic, positions read from the step's start."""
    _lib.oracle_step_snapshot_f32(_fp(X), _fp(A), _fp(V), len(X), C.c_float(eps2), C.c_float(dt))


def accel_f32(X, idx, eps2=0.002):
    idx = np.ascontiguousarray(idx, np.int32)
    out = np.zeros((len(idx), 3), np.float32)
    _lib.oracle_accel_f32_sampled(_fp(X), len(X), C.c_float(eps2), idx.ctypes.data_as(_I), len(idx), _fp(out))
    return out


def accel_f64(X, idx=None, eps2=0.002):
    if idx is None:
        idx = np.arange(len(X), dtype=np.int32)
    idx = np.ascontiguousarray(idx, np.int32)
    out = np.zeros((len(idx), 3), np.float64)
    _lib.oracle_accel_f64_sampled(_fp(X), len(X), C.c_double(eps2), idx.ctypes.data_as(_I), len(idx), out.ctypes.data_as(_D))
    return out


def energy(X, V, eps2=0.002):
    """(kinetic, pair potential sum_{i<j}); the scheme conserves ke + 0.5*pe"""
    out = np.zeros(2, np.float64)
    _lib.oracle_energy_f64(_fp(X), _fp(V), len(X), C.c_double(eps2), out.ctypes.data_as(_D))
    return float(out[0]), float(out[1])


def step_headless_f3(X, V3, eps2=0.002, dt=0.01):
    _lib.oracle_step_headless_f3(_fp(X), _fp(V3), len(X), C.c_double(eps2), C.c_double(dt))


def ic_reference(n, max_xyz=1e5, min_w=1e5, max_w=1e9):
    X = np.zeros((n, 4), np.float32)
    _lib.oracle_ic_reference(_fp(X), n, C.c_float(max_xyz), C.c_float(min_w), C.c_float(max_w))
    return X


def verify_still_bodies(v, x) -> int:
    _lib.oracle_verify_still_bodies.restype = C.c_int
    return int(_lib.oracle_verify_still_bodies(_fp(v), _fp(x), len(x)))


# ---- the reference's own object code (present only when oracle/_ref was built) ----------
def ref_lib(omp=False):
    """omp: False = serial build (as the vcxproj), True = -O2 -fopenmp, "o3" = -O3 -march=x86-64-v3 -fopenmp"""
    p = os.path.join(HERE, "_ref", "libref_cpu_omp_o3.so" if omp == "o3" else "libref_cpu_omp.so" if omp else "libref_cpu.so")
    return C.CDLL(p) if os.path.exists(p) else None


def ref_cpu_compute(X, A, V, omp=False):
    """reference CPU_compute, literal (in place).  Constants are the reference's own (EPS2 0.002f, DT 0.1f)."""
    r = ref_lib(omp)
    r.ref_CPU_compute(_fp(X), _fp(A), _fp(V), len(X))


def ref_snapshot_step(X, A, V, omp=False):
    r = ref_lib(omp)
    r.ref_snapshot_step(_fp(X), _fp(A), _fp(V), len(X))


def ref_fill_with_random4(n):
    X = np.zeros((n, 4), np.float32)
    ref_lib().ref_fill_with_random4(_fp(X), n)
    return X


def ref_pair(bi, bj, ai):
    ai = np.array(ai, np.float32)
    ref_lib().ref_bodyInteractions_CPU(_fp(np.asarray(bi, np.float32)), _fp(np.asarray(bj, np.float32)), _fp(ai))
    return ai


# ---- error measures used by every parity test ---------------------------------------------------
def norm_rel_err(a, b):
    """max over bodies of |a_i - b_i| / |b_i| on the xyz components (per-body
    norm-relative, never per-component)"""
    a = np.asarray(a, np.float64)[:, :3]
    b = np.asarray(b, np.float64)[:, :3]
    den = np.linalg.norm(b, axis=1)
    den = np.where(den > 0, den, 1.0)
    return float(np.max(np.linalg.norm(a - b, axis=1) / den))


def step_mixed(X, A, V, eps2=0.002, dt=0.1):
    """reduction flavours (kernel.cu:406-436 / :439-469): double half-kick, float FMA drift.
    V, A are (N,4) or (N,3) float32.  Pinned on the reference kernels' output on a B200
    (tests/golden/headless_ref_gpu.npz)."""
    stride = V.shape[1]
    assert A.shape[1] == stride and stride in (3, 4)
    _lib.oracle_step_mixed(_fp(X), _fp(A), _fp(V), len(X), C.c_float(eps2), C.c_float(dt), C.c_int(stride))


def render(X, width=800, height=600, max_xyz=(2e5, 2e5, 2e5), min_w=1e5, max_w=1e9, depth_test=True):
    """point-splat frame of simulation_visualization.cpp's shaders; returns uint8 (H, W, 4), top row first"""
    img = np.zeros((height, width, 4), np.uint8)
    _lib.oracle_render(_fp(X), len(X), C.c_int(width), C.c_int(height), C.c_float(max_xyz[0]), C.c_float(max_xyz[1]),
                       C.c_float(max_xyz[2]), C.c_float(min_w), C.c_float(max_w), C.c_int(1 if depth_test else 0),
                       img.ctypes.data_as(C.c_void_p))
    return img
