"""Runs the headless driver binary (n-bodysimulation_b200/nbody_headless, csrc/nbody_headless.cpp), the restatement
of the reference's own headless run: TestProject/main.cpp:142-160 (simulationLoopNoVisual: `steps` calls of
simulate(), "Starting the simulation..." / "Simulation complete", exceptions -> EXIT_FAILURE), :244-284 (pinned host
state, initial conditions, device copies) and :350-359.  Its printed state checksums are compared with the checker's
state for every flavour of the launch interface, its resume path with a straight run, its multi-GPU forms with the
one-GPU run."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "n-bodysimulation_b200", "nbody_headless")


def headless(*args, ok=True):
    r = subprocess.run([EXE] + [str(a) for a in args], capture_output=True, text=True, timeout=300)
    if ok:
        assert r.returncode == 0, (r.returncode, r.stdout[-400:], r.stderr[-400:])
    return r


def result(r):
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, r.stdout
    return json.loads(lines[0])


def checksum(a):
    """the driver's checksum (nbody_headless.cpp): sum of x + 2 y + 3 z in double, and its scale"""
    a = np.asarray(a, np.float64)
    return float((a[:, 0] + 2 * a[:, 1] + 3 * a[:, 2]).sum()), float((np.abs(a[:, 0]) + 2 * np.abs(a[:, 1]) + 3 * np.abs(a[:, 2])).sum())


def assert_checksums(res, x, v, tol_x=1e-6, tol_v=1e-5):
    # per-body errors are <= 1e-6 (x) / 1e-5 (v) norm-relative (tests/test_parity_gpu.py); a sum of N of them is bounded
    # by the same fraction of the sum of magnitudes
    cx, sx = checksum(x)
    cv, sv = checksum(v)
    assert abs(res["checksum_x"] - cx) <= tol_x * sx, (res["checksum_x"], cx, sx)
    assert abs(res["checksum_v"] - cv) <= tol_v * max(sv, 1e-30), (res["checksum_v"], cv, sv)


def test_binary_exists():
    assert os.access(EXE, os.X_OK), "n-bodysimulation_b200/nbody_headless not built (python -c 'import __graft_entry__ as g; g.build()')"


def test_reference_run_config1_against_the_reference_golden_state(cuda, golden):
    # BASELINE configs[0]: N = 1024, the reference's own initial conditions (fill_with_random4 under rand()'s default
    # seed), 10 steps; tests/golden/steps_n1024.npz was generated from the reference's object code
    g = golden["steps_n1024"]
    steps = int(g["steps"])
    r = headless("--n", 1024, "--steps", steps, "--ic", "reference", "--json")
    out = r.stdout.splitlines()
    assert out[0] == "Starting the simulation..." and out[1] == "Simulation complete"      # main.cpp:144,158
    res = result(r)
    assert (res["n"], res["steps"], res["flavour"], res["gpus"]) == (1024, steps, "simulate", 1)
    assert res["launches"] == 2 * steps                                                    # pack + step kernel per simulate()
    assert_checksums(res, g["snapshot_x"], g["snapshot_v"], tol_x=1e-5, tol_v=1e-4)      # 10 steps: the bounds of test_config1_n1024_10_steps_vs_reference_golden


@pytest.mark.parametrize("flavour", ["simulate", "emb", "emb-float3"])
def test_every_launcher_flavour_against_the_checker(cuda, nb, oracle, flavour):
    n, steps = 3000, 3
    res = result(headless("--n", n, "--steps", steps, "--ic", "uniform", "--seed", 5, "--flavour", flavour, "--json"))
    assert res["flavour"] == flavour
    X = nb.ic_uniform(n, seed=5)
    if flavour == "emb-float3":
        V, A = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
    else:
        V, A = np.zeros((n, 4), np.float32), np.zeros((n, 4), np.float32)
    for _ in range(steps):
        if flavour == "simulate":
            oracle.step_snapshot(X, A, V)
        else:
            oracle.step_mixed(X, A, V)
    if flavour == "emb-float3":
        # the driver views its float4 allocations as packed float3 arrays and prints the checksum over the float4 view:
        # compare positions only (the velocity checksum of that view mixes components)
        cx, sx = checksum(X)
        assert abs(res["checksum_x"] - cx) <= 1e-6 * sx
    else:
        assert_checksums(res, X, V)


def test_host_pointer_flavour(cuda, nb, oracle):
    n, steps = 2048, 2
    res = result(headless("--n", n, "--steps", steps, "--ic", "uniform", "--seed", 6, "--host-f3", "--json"))
    assert res["flavour"] == "host-f3"
    X = nb.ic_uniform(n, seed=6)
    V3 = np.zeros((n, 3), np.float32)
    for _ in range(steps):
        oracle.step_headless_f3(X, V3)
    assert_checksums(res, X, V3)


def test_resident_save_and_load_resume_equals_the_straight_run(cuda, tmp_path):
    snap = str(tmp_path / "run.nbody")
    common = ("--n", 5000, "--ic", "plummer", "--seed", 3, "--dt", 0.01, "--resident", "--json")
    straight = result(headless(*common, "--steps", 9))
    first = result(headless(*common, "--steps", 4, "--save", snap))
    assert first["flavour"] == "resident" and os.path.getsize(snap) == 64 + 2 * 16 * 5000
    resumed = result(headless(*common, "--steps", 5, "--load", snap))
    assert resumed["checksum_x"] == straight["checksum_x"] and resumed["checksum_v"] == straight["checksum_v"]   # bit-identical
    # the reference-shaped loop from the same initial conditions lands on the same state (same kernels, same order)
    loop = result(headless("--n", 5000, "--ic", "plummer", "--seed", 3, "--dt", 0.01, "--steps", 9, "--json"))
    assert loop["checksum_x"] == straight["checksum_x"]


def test_failures_exit_like_the_reference(cuda, tmp_path):
    # simulate() throws -> caught as std::exception, message on stderr, EXIT_FAILURE (main.cpp:150-153)
    r = headless("--n", 1000, "--steps", 1, "--eps2", 0, ok=False)
    assert r.returncode == 1 and "eps2" in r.stderr and "Simulation complete" not in r.stdout
    r = headless("--n", 0, ok=False)
    assert r.returncode == 2
    r = headless("--n", 100, "--resident", "--load", str(tmp_path / "missing.nbody"), ok=False)
    assert r.returncode == 1 and "cannot open" in r.stderr
    r = headless("--n", 100, "--gpus", 99, ok=False)
    assert r.returncode == 2
    r = headless("--frobnicate", ok=False)
    assert r.returncode == 2


def test_frame_of_the_last_step(cuda, tmp_path):
    ppm = str(tmp_path / "f.ppm")
    headless("--n", 4096, "--steps", 1, "--ic", "reference", "--ppm", ppm)
    head = open(ppm, "rb").read(15)
    assert head.startswith(b"P6\n800 600\n255\n") and os.path.getsize(ppm) == 15 + 800 * 600 * 3


@pytest.mark.parametrize("mode", ["resident-p2p", "resident-nccl", "set-devices"])
def test_two_gpus_from_one_process(cuda, mode):
    if cuda.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    # Plummer sphere, dt = 0.01: a smooth system (the uniform cube at DT = 0.1 has close pairs that fly 5 000 units
    # per step from a separation of 30 and amplify any rounding difference a thousandfold in four steps)
    n, steps = 100000, 4
    common = ("--n", n, "--steps", steps, "--ic", "plummer", "--seed", 7, "--dt", 0.01, "--json")
    one = result(headless(*common, "--resident"))
    if mode == "set-devices":
        two = result(headless(*common, "--gpus", 2))               # the reference's loop, simulate_set_devices(2)
        assert two["exchange"] == "set_devices"
    else:
        two = result(headless(*common, "--gpus", 2, "--resident", "--exchange", mode.split("-")[1]))
        assert two["exchange"] == mode.split("-")[1]
    assert two["gpus"] == 2
    # other summation order: checksums agree to rounding (|x| ~ 1, |v| ~ 0.5: sums of magnitudes ~ 3 N)
    assert abs(two["checksum_x"] - one["checksum_x"]) <= 1e-6 * 3 * n
    assert abs(two["checksum_v"] - one["checksum_v"]) <= 1e-5 * 3 * n


@pytest.mark.parametrize("n", [12288, 16384, 20000])
def test_spread_fixup_of_split_tiles_same_bits_as_the_last_arriver(cuda, nb, n):
    # mid-size N: a split tile is finished by all of its CTAs (force_kernel's SPREAD, cooperative launch) instead of the
    # last one to arrive: the same partial sums added in the same order -> identical state, in the resident (graph
    # replayed) loop and in the reference-shaped simulate() loop
    # (NBODY_VARIANT_AUTO without the row-split family, which takes 16 384 since round 2 and splits no tile at all)
    with nb.Simulation(n, variant=110) as sim:
        assert sim.info()["spread_fixup"]
    with nb.Simulation(1 << 17) as sim:
        assert not sim.info()["spread_fixup"]                     # large N: the plain launch
    for fl in (("--resident",), ()):
        common = ("--n", n, "--steps", 3, "--ic", "plummer", "--seed", 4, "--dt", 0.01, "--json") + fl
        env = dict(os.environ, NBODY_NO_ROWSPLIT="1")
        r = subprocess.run([EXE] + [str(x) for x in common], capture_output=True, text=True, timeout=300, env=env)
        assert r.returncode == 0, r.stderr[-400:]
        a = result(r)
        env = dict(os.environ, NBODY_NO_ROWSPLIT="1", NBODY_NO_SPREAD_FIXUP="1")
        r = subprocess.run([EXE] + [str(x) for x in common], capture_output=True, text=True, timeout=300, env=env)
        assert r.returncode == 0, r.stderr[-400:]
        b = result(r)
        assert a["checksum_x"] == b["checksum_x"] and a["checksum_v"] == b["checksum_v"]


def test_energy_lines_of_a_resident_run(cuda):
    # BASELINE's drift bound through the driver: Plummer sphere, dt = 0.01, |E_eff drift| <= 1e-3 over 100 steps
    # (tests/test_parity_gpu.py holds the same bound at config 3 through the C ABI)
    r = headless("--n", 20000, "--steps", 100, "--ic", "plummer", "--seed", 9, "--dt", 0.01, "--resident", "--energy", "--json")
    lines = [json.loads(ln) for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 2 and "rel_drift" in lines[1]
    e = lines[1]
    assert e["ke0"] > 0 and e["pe0"] < 0 and abs(e["rel_drift"]) <= 1e-3, e
    assert headless("--n", 100, "--energy", ok=False).returncode == 2      # needs --resident
