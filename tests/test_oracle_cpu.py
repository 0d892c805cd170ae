"""CPU tests of the checker itself: oracle/nbody_oracle.c against the golden
vectors made from the reference's object code (tests/golden/make_golden.py) and,
where oracle/_ref exists, against that object code directly."""
import numpy as np
import pytest


def test_pair_matches_reference_golden(oracle, golden):
    g = golden["pairs"]
    for k in range(len(g["bi"])):
        got = oracle.pair(g["bi"][k], g["bj"][k], g["ai"][k])
        assert np.array_equal(got[:3], g["out"][k][:3]), k   # bit exact (validation.cpp:9-24)


def test_ic_reference_matches_golden(oracle, golden):
    assert np.array_equal(oracle.ic_reference(1024), golden["ic_reference"]["x1024"])   # utils.cpp:30-37


@pytest.mark.parametrize("n", [256, 1024])
def test_steps_match_reference_golden(oracle, golden, n):
    g = golden[f"steps_n{n}"]
    for name, fn in (("literal", oracle.step_literal), ("snapshot", oracle.step_snapshot)):
        X, V, A = g["x0"].copy(), np.zeros_like(g["x0"]), np.zeros_like(g["x0"])
        fn(X, A, V)
        assert np.array_equal(A, g[f"{name}_a1"])
        for _ in range(int(g["steps"]) - 1):
            fn(X, A, V)
        assert np.array_equal(X, g[f"{name}_x"]) and np.array_equal(V, g[f"{name}_v"]) and np.array_equal(A, g[f"{name}_a"])


def test_against_reference_object_code_when_present(oracle):
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built on this machine (reference sources absent)")
    n = 300   # ragged size, not in the golden set
    X0 = oracle.ref_fill_with_random4(n)
    for ours, theirs in ((oracle.step_literal, oracle.ref_cpu_compute), (oracle.step_snapshot, oracle.ref_snapshot_step)):
        Xa, Va, Aa = X0.copy(), np.zeros_like(X0), np.zeros_like(X0)
        Xb, Vb, Ab = X0.copy(), np.zeros_like(X0), np.zeros_like(X0)
        for _ in range(3):
            ours(Xa, Aa, Va)
            theirs(Xb, Ab, Vb)
        assert np.array_equal(Xa, Xb) and np.array_equal(Va, Vb) and np.array_equal(Aa, Ab)


def test_literal_vs_snapshot_is_the_references_own_inconsistency(oracle, golden):
    # in-place CPU_compute differs from snapshot semantics at the 1e-3 level after 1 step
    g = golden["steps_n1024"]
    e = oracle.norm_rel_err(g["literal_a1"], g["snapshot_a1"])
    assert 1e-5 < e < 1e-1


def test_fp32_snapshot_close_to_fp64(oracle, golden):
    g = golden["steps_n1024"]
    a64 = oracle.accel_f64(g["x0"])
    assert oracle.norm_rel_err(g["snapshot_a1"], a64) < 1e-5
    idx = np.array([0, 5, 1023], np.int32)
    assert np.array_equal(oracle.accel_f32(g["x0"], idx), g["snapshot_a1"][idx, :3])


def test_empty_and_tiny(oracle):
    X = np.array([[1, 2, 3, 5]], np.float32)
    V, A = np.zeros_like(X), np.ones_like(X)
    oracle.step_snapshot(X, A, V)
    assert np.array_equal(A, np.zeros_like(A)) and np.array_equal(X, [[1, 2, 3, 5]])   # one body: no force
    X = np.array([[0, 0, 0, 1], [1, 0, 0, 1]], np.float32)
    V, A = np.zeros_like(X), np.zeros_like(X)
    oracle.step_snapshot(X, A, V, eps2=0.002, dt=0.1)
    exp = np.float32(1.0) / np.sqrt(np.float32(1.002) ** 3, dtype=np.float32)
    assert np.isclose(A[0, 0], exp, rtol=1e-6) and A[1, 0] == -A[0, 0]


def test_energy_is_conserved_by_the_scheme_in_eff_form(oracle):
    # half-kick-only scheme conserves KE + 0.5*PE, not KE + PE
    rng = np.random.default_rng(1)
    n = 64
    X = np.concatenate([rng.normal(0, 1, (n, 3)), np.full((n, 1), 1.0 / n)], 1).astype(np.float32)
    V, A = np.zeros_like(X), np.zeros_like(X)
    ke0, pe0 = oracle.energy(X, V, 0.01)
    for _ in range(200):
        oracle.step_snapshot(X, A, V, eps2=0.01, dt=0.01)
    ke1, pe1 = oracle.energy(X, V, 0.01)
    eff = abs((ke1 + 0.5 * pe1) - (ke0 + 0.5 * pe0)) / abs(ke0 + 0.5 * pe0)
    full = abs((ke1 + pe1) - (ke0 + pe0)) / abs(ke0 + pe0)
    assert eff < 5e-3 and eff < 0.1 * full


# The flavours the reference has as GPU code only are pinned on that code's output on a B200
# (tests/golden/make_golden_headless_gpu.py: the reference's kernels rebuilt for sm_100 by oracle/Makefile).  The
# restatement sums in index order without FMA contraction, the GPU kernels in their own order (tiles of 1024 in
# sequence; 32-wide trees, then blocks in sequence) with FMA: agreement is to summation rounding -- measured
# 5.6e-7 .. 7.6e-7 (headless flavour) and 1.4e-6 .. 4.4e-6 (reduction flavours, masses up to 1e9 cancel harder).
# Positions are stored as exact differences from the initial ones (see slim() in the generator).
@pytest.mark.parametrize("n", [1024, 4096])
def test_headless_flavour_matches_the_reference_kernel_output(oracle, golden, n):
    g = golden["headless_ref_gpu"]                       # Sim-Without-OpenGL-Integration/kernel.cu:38-125 on a B200
    x0 = oracle.ic_reference(n, max_w=1e8)               # headless constants.h: MAX_W 1e8
    X, V3 = x0.copy(), np.zeros((n, 3), np.float32)
    oracle.step_headless_f3(X, V3, 0.002, 0.01)
    assert oracle.norm_rel_err(V3, g[f"v1_n{n}"]) <= 2e-6 and np.array_equal(X, x0 + g[f"d_x1_n{n}"])
    for _ in range(int(g["steps"]) - 1):
        oracle.step_headless_f3(X, V3, 0.002, 0.01)
    assert oracle.norm_rel_err(V3, g[f"vk_n{n}"]) <= 2e-6 and oracle.norm_rel_err(X, x0 + g[f"d_xk_n{n}"]) <= 1e-7


@pytest.mark.parametrize("stride", [4, 3])
@pytest.mark.parametrize("n", [1024, 2048])
def test_reduction_flavour_matches_the_reference_kernel_output(oracle, golden, n, stride):
    g = golden["headless_ref_gpu"]                       # kernel.cu:163-318 + :406-469 on a B200
    assert int(g["red3_equals_red4_xyz"]) == 1           # the float3 twin gave the same xyz bits; stored once
    x0 = oracle.ic_reference(n)
    X, V, A = x0.copy(), np.zeros((n, stride), np.float32), np.zeros((n, stride), np.float32)
    oracle.step_mixed(X, A, V)
    k = f"red4_%s_n{n}"
    assert oracle.norm_rel_err(A, g[k % "a1"]) <= 1e-5 and oracle.norm_rel_err(V, g[k % "v1"]) <= 1e-5
    assert oracle.norm_rel_err(X, x0 + g["d_" + k % "x1"]) <= 1e-7
    for _ in range(int(g["steps"]) - 1):
        oracle.step_mixed(X, A, V)
    assert oracle.norm_rel_err(V, g[k % "vk"]) <= 1e-5 and oracle.norm_rel_err(X, x0 + g["d_" + k % "xk"]) <= 1e-7


def test_headless_flavour_tracks_main_flavour(oracle):
    # the two flavours share the pair sum and differ by the double-vs-float integrator rounding only
    X0 = oracle.ic_reference(200)
    Xa, Va, Aa = X0.copy(), np.zeros_like(X0), np.zeros_like(X0)
    oracle.step_snapshot(Xa, Aa, Va, eps2=0.002, dt=0.01)
    Xb, Vb = X0.copy(), np.zeros((200, 3), np.float32)
    oracle.step_headless_f3(Xb, Vb, 0.002, 0.01)
    assert oracle.norm_rel_err(Xb, Xa) < 1e-6 and oracle.norm_rel_err(Vb, Va) < 1e-5


def test_verify_still_bodies_restatement(oracle):
    a = np.array([[100, 100, 100, 0], [100, 100, 100, 0]], np.float32)
    b = np.array([[100.5, 100, 100, 0], [102, 100, 100, 0]], np.float32)
    assert oracle.verify_still_bodies(a, b) == 1   # validation.cpp:143-164: 1 % per component
