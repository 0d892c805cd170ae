"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol
include/nbody_b200.h declares plus the reference's mangled simulate() overloads,
validates arguments, and refuses to compute without a CUDA device (no fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "nbody_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nbody_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(nb):
    names = declared_symbols()
    assert len(names) >= 20
    for s in names:
        assert hasattr(nb.lib, s), f"{s} declared in include/nbody_b200.h but not exported"
    assert set(names) == set(nb.SIGNATURES), set(names) ^ set(nb.SIGNATURES)


def test_library_exports_reference_launch_interface(nb):
    out = subprocess.run(["nm", "-D", "--defined-only", nb.LIB_PATH], capture_output=True, text=True, check=True).stdout
    assert "_Z8simulateP6float4S0_S0_i" in out          # TestProject/kernel.cuh:2
    assert "_Z8simulateP6float4P6float3S2_i" in out     # Sim-Without-OpenGL-Integration/kernel.cuh:5


def test_params_default_are_the_reference_constants(nb):
    p = nb.NBodyParams()
    assert nb.lib.nbody_params_default(C.byref(p)) == 0
    assert (p.n_bodies, p.world, p.rank) == (8192, 1, 0)                 # constants.h:13
    assert p.eps2 == np.float32(0.002) and p.dt == np.float32(0.1)       # constants.h:25-26
    assert (p.eps2_f64, p.dt_f64) == (0.002, 0.01)                       # headless constants.h:14-15


def test_argument_validation(nb):
    h = C.c_void_p()
    p = nb.NBodyParams()
    nb.lib.nbody_params_default(C.byref(p))
    for field, bad in (("n_bodies", 0), ("n_bodies", -5), ("eps2", 0.0), ("world", 0), ("rank", 3), ("integrator", 7), ("summation", 7)):
        q = nb.NBodyParams.from_buffer_copy(p)
        setattr(q, field, bad)
        assert nb.lib.nbody_create(C.byref(h), C.byref(q)) == -1, field
        assert nb.lib.nbody_last_error()
    # eps2 whose cube is subnormal: MUFU.RSQ would flush it and turn the self term into 0 * inf (the reference's
    # 1/sqrtf stays finite there) -- rejected, not answered with NaNs; the bound itself (2^-42) is accepted up to
    # the point where a device is needed
    q = nb.NBodyParams.from_buffer_copy(p)
    q.eps2 = 1e-13
    assert nb.lib.nbody_create(C.byref(h), C.byref(q)) == -1 and b"too small" in nb.lib.nbody_last_error()
    assert nb.lib.nbody_step_inplace_f4(1, 1, 1, 16, 1e-13, 0.1) == -1
    q.eps2 = 2.0 ** -42
    assert nb.lib.nbody_create(C.byref(h), C.byref(q)) in (0, -2)   # valid argument: succeeds on a GPU, NBODY_E_CUDA without
    if h.value:
        nb.lib.nbody_destroy(h)
    assert nb.lib.nbody_step_inplace_f4(None, None, None, 16, 0.002, 0.1) == -1
    assert nb.lib.nbody_step_inplace_f4(1, 1, 1, 0, 0.002, 0.1) == -1
    assert nb.lib.nbody_step_host_f3(None, None, None, 4, 0.002, 0.01) == -1
    f = C.c_int(7)
    assert nb.lib.nbody_ctx_fixup(None, C.byref(f)) == -1 and f.value == 7 and b"null" in nb.lib.nbody_last_error()


def test_multi_device_parameters_are_validated_on_the_host(nb):
    # nbody_params.n_devices > 1: ONE context over several GPUs ("nbody_create takes a device list")
    h = C.c_void_p()
    p = nb.NBodyParams()
    nb.lib.nbody_params_default(C.byref(p))
    assert (p.n_devices, p.exchange, p.deterministic, p.summation) == (0, nb.NBODY_EXCHANGE_P2P, 0, nb.NBODY_SUM_TWO_LEVEL)
    q = nb.NBodyParams.from_buffer_copy(p)
    q.n_devices = nb.NBODY_MAX_DEVICES + 1
    assert nb.lib.nbody_create(C.byref(h), C.byref(q)) == -1
    q.n_devices, q.world, q.rank = 2, 2, 1                     # a group is the whole system
    assert nb.lib.nbody_create(C.byref(h), C.byref(q)) == -1 and b"whole system" in nb.lib.nbody_last_error()
    q.world, q.rank, q.exchange = 1, 0, 7
    assert nb.lib.nbody_create(C.byref(h), C.byref(q)) == -1 and b"exchange" in nb.lib.nbody_last_error()
    assert nb.lib.nbody_group_info(None, None, None, None) == -1
    assert nb.lib.nbody_shard_barrier_wait_ns(None, None, None, None, 0) == -1
    # nbody_set_devices: list checks need the device count, so without a GPU the call reports NBODY_E_CUDA; clearing works anywhere
    assert nb.lib.nbody_set_devices(None, 0) == 0
    arr = (C.c_int * 17)(*range(17))
    assert nb.lib.nbody_set_devices(arr, 17) == -1


def test_no_cpu_fallback(nb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("has a GPU")
    with pytest.raises(nb.NBodyError):
        nb.Simulation(1024)
    x = np.zeros((8, 4), np.float32)
    with pytest.raises(nb.NBodyError):
        nb.simulate_host_f3(x, np.zeros((8, 3), np.float32), np.zeros((8, 3), np.float32), 8)


def test_shard_geometry(nb):
    f = nb.lib.nbody_shard_slot_bytes
    assert f(1 << 20, 8) == (1 << 17) * 16
    assert f(1000, 3) == 1024 * 16         # per = 334 bodies, padded to the 1024-body slot grain
    assert f(1025, 2) == 1024 * 16         # per = 514 (even ceil) -> 1024
    assert f(0, 1) == 0 and f(10, 0) == 0


def test_variants_listed(nb):
    v = nb.variants()
    assert len(v) >= 8 and all("targets/thread" in s for s in v.values())


def test_ic_uniform_is_seeded_and_in_range(nb):
    a, b, c = nb.ic_uniform(4096, seed=1), nb.ic_uniform(4096, seed=1), nb.ic_uniform(4096, seed=2)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert np.abs(a[:, :3]).max() <= 1e5 and a[:, 3].min() >= 1e5 and a[:, 3].max() <= 1e9
    assert abs(a[:, :3].mean()) < 3e3 and abs(a[:, 3].mean() - 5e8) < 2e7
    assert np.array_equal(nb.ic_uniform(100, seed=1), a[:100])   # counter-based: prefix-stable


def test_ic_plummer_is_a_plummer_sphere(nb, oracle):
    x, v = nb.ic_plummer(4096, seed=3)
    assert np.isclose(x[:, 3].sum(), 1.0, rtol=1e-5)
    assert np.abs(x[:, :3].mean(0)).max() < 1e-6 and np.abs(v[:, :3].mean(0)).max() < 1e-6
    r = np.linalg.norm(x[:, :3], axis=1)
    assert 0.6 < np.median(r) / (1.3048 * 3 * np.pi / 16) < 1.4    # half-mass radius = 1.3048 a
    ke, pe = oracle.energy(x, v, 1e-6)
    # virial equilibrium for the scheme's effective G = 1/2: 2*KE = -0.5*PE
    assert 0.8 < 2 * ke / (-0.5 * pe) < 1.2


def test_kernel_choice_rules_are_host_logic(nb):
    # NBODY_VARIANT_AUTO by size (DESIGN.md section 3 "Variant choice" / "Mid-size N"); host-only, no GPU involved
    names = nb.variants()
    pick = lambda t, s=None: nb.choose_variant(t, s or t)   # noqa: E731
    assert names[pick(1024)[0]].startswith("S_TT2") and names[pick(2300)[0]].startswith("S_TT2")   # small-N family
    assert names[pick(3200)[0]].startswith("S_TT4") and names[pick(3000)[0]].startswith("R_T6_H4")
    assert pick(1024) == (85, False) and pick(2047) == (85, False)
    # unsharded mid sizes whose whole target tiles fill 148 SMs evenly: row-split family (tile = T x H targets per CTA,
    # one wave); 16 384 = BASELINE config 2 -> 147 tiles of 7 x 16
    assert names[pick(16384)[0]].startswith("R_T7_H16") and names[pick(8192)[0]].startswith("R_T7_H8")
    assert names[pick(4096)[0]].startswith("R_T7_H4") and names[pick(18944)[0]].startswith("R_T8_H16")
    assert names[pick(10240)[0]].startswith("R_T9_H8") and names[pick(20480)[0]].startswith("R_T9_H16")
    # ... otherwise the stream-K kernels, tile shape x unit size by unit balance (DESIGN.md section 3 "Mid-size N")
    assert names[pick(22000)[0]].startswith("T8_NT256_UJ128") and names[pick(50000)[0]].startswith("T6_NT256_UJ256")
    assert all(not pick(n)[1] and not names[pick(n)[0]].startswith("S_") for n in (8192, 10240, 12288, 22000, 65535))
    assert all(names[pick(n)[0]].startswith("T") for n in (12288, 17500, 22000, 26000, 40000, 65535))
    # a shard never gets the row-split family (it has no in-kernel barrier)
    assert names[pick(16384, 131072)[0]].startswith("T")
    # large: one of the equivalent kernels, timed on the device
    assert pick(65536)[1] and pick(1 << 20)[1] and pick(131072, 1 << 20)[1]
    # a shard with few targets of many sources: more CTAs per SM instead of more targets per thread
    assert names[pick(4096, 1 << 20)[0]].startswith("T2_NT256") and names[pick(1000, 1 << 20)[0]].startswith("T1_NT128")
    with pytest.raises(nb.NBodyError):
        nb.choose_variant(0, 10)
    with pytest.raises(nb.NBodyError):
        nb.choose_variant(20, 10)


def test_row_split_rule_only_takes_whole_single_waves(nb):
    # the row-split family runs one CTA per tile of T x H targets: NBODY_VARIANT_AUTO may take it only where the tiles
    # fit one wave of the 148 SMs and fill most of it (measured never slower there, profiles/r2_rowsplit.txt), and
    # never for a shard (it has no in-kernel barrier)
    names = nb.variants()
    picked = 0
    for n in range(2048, 65536, 61):
        name = names[nb.choose_variant(n, n)[0]]
        m = re.match(r"R_T(\d+)_H(\d+)_", name)
        if not m:
            continue
        picked += 1
        tile = int(m.group(1)) * int(m.group(2))
        assert -(-n // tile) <= 148, (n, name)
        assert n / (148.0 * tile) >= (0.85 if n >= 8192 else 0.68), (n, name)
    assert picked > 100
    assert names[nb.choose_variant(40000, 40000)[0]].startswith("T") and names[nb.choose_variant(65535, 65535)[0]].startswith("T")
    for targets in (4096, 8192, 16384, 32768):
        for sources in (65536, 131072, 1 << 20):
            assert not names[nb.choose_variant(targets, sources)[0]].startswith("R_")
