"""Row-split family (rowsplit_kernel, csrc/nbody_kernels.cuh): whole target tiles per CTA, every CTA streams all
sources, the source split reduced inside the CTA.  NBODY_VARIANT_AUTO picks it for unsharded mid-size systems whose
tiles fill the SMs evenly -- BASELINE config 2 (N = 16 384) among them.  Checked against the oracle
(validation.cpp:9-52 restated, snapshot reading) with the tolerances of tests/test_parity_gpu.py, and against the
stream-K kernel on the same input."""
import numpy as np
import pytest

TOL_A, TOL_X, TOL_V = 1e-5, 1e-6, 1e-5
pytestmark = pytest.mark.gpu


def one_step(nb, x0, v0=None, **kw):
    with nb.Simulation(len(x0), **kw) as sim:
        sim.upload(x0, v0).step(1)
        return sim.download(), sim.info()


# 4 144 = 148 x 28 (4 lanes x T7), 8 192 (8 x T7; last unit full), 8 325 (ragged tile, ragged last unit, odd N),
# 11 840 = 148 x 80, 16 384 = config 2 (147 tiles of 112), 18 944 = 148 x 128, 23 680 = 148 x 160 (32 lanes x T5)
@pytest.mark.parametrize("n", [4144, 8192, 8325, 11840, 16384, 18944, 23680])
def test_auto_picks_row_split_and_matches_oracle(cuda, nb, oracle, n):
    x0 = nb.ic_uniform(n, seed=n)
    v0 = np.zeros_like(x0)
    v0[:, :3] = np.random.default_rng(n).standard_normal((n, 3)).astype(np.float32)
    (x, v, a), info = one_step(nb, x0, v0)
    assert info["variant_name"].startswith("R_T"), info
    a64 = oracle.accel_f64(x0)
    X, V, A = x0.copy(), v0.copy(), np.zeros_like(x0)
    oracle.step_snapshot(X, A, V)
    ref_err = oracle.norm_rel_err(A, a64)
    err = oracle.norm_rel_err(a, a64)
    print(f"N={n} {info['variant_name']}: err(ours, fp64)={err:.2e} err(O2, fp64)={ref_err:.2e} err(ours, O2)={oracle.norm_rel_err(a, A):.2e}")
    assert err <= max(TOL_A, 1.5 * ref_err)                    # as config 2 in tests/test_parity_gpu.py
    assert oracle.norm_rel_err(a, A) <= max(3 * TOL_A, err + 1.05 * ref_err)   # O2 itself is 3.4e-5 from fp64 at N = 23 680
    assert oracle.norm_rel_err(x, X) <= TOL_X and oracle.norm_rel_err(v, V) <= 3 * TOL_V
    assert np.array_equal(x[:, 3], x0[:, 3]) and np.all(a[:, 3] == 0)     # mass kept, A.w = 0 (kernel.cu:127-129)


def test_row_split_agrees_with_stream_k_and_is_deterministic(cuda, nb, oracle):
    n = 16384
    x0 = nb.ic_uniform(n, seed=5)
    (x, v, a), info = one_step(nb, x0)
    (x2, v2, a2), _ = one_step(nb, x0)
    assert np.array_equal(a, a2) and np.array_equal(x, x2)     # fixed reduction order
    (xs, vs, as_), info_s = one_step(nb, x0, variant=110)      # the stream-K kernel AUTO used before
    assert info["variant_name"].startswith("R_T7_H16") and info_s["variant_name"].startswith("T8_")
    assert oracle.norm_rel_err(a, as_) <= 2 * TOL_A and oracle.norm_rel_err(x, xs) <= TOL_X


def test_row_split_several_steps_track_the_stream_k_run(cuda, nb, oracle):
    # the packed next-buffer writes of the row-split epilogue feed the following steps
    n = 8192
    x0, v0 = nb.ic_plummer(n, seed=3)
    out = []
    for variant in (0, 116):
        with nb.Simulation(n, eps2=1e-4, dt=1e-3, variant=variant) as sim:
            sim.upload(x0, v0).step(40)
            out.append(sim.download())
            name = sim.info()["variant_name"]
        assert name.startswith("R_T") == (variant == 0)
    assert oracle.norm_rel_err(out[0][0], out[1][0]) <= 1e-5 and oracle.norm_rel_err(out[0][1], out[1][1]) <= 1e-4


def test_in_place_simulate_uses_it_too(cuda, nb, oracle):
    torch = cuda
    n = 16384
    x0 = nb.ic_uniform(n, seed=9)
    dx, dv, da = torch.from_numpy(x0).cuda(), torch.zeros(n, 4, device="cuda"), torch.zeros(n, 4, device="cuda")
    nb.simulate(dx, da, dv, n)                                  # kernel.cuh:2
    (x, v, a), _ = one_step(nb, x0)
    assert np.array_equal(da.cpu().numpy(), a) and np.array_equal(dx.cpu().numpy(), x)


def test_snapshots_of_a_row_split_run(cuda, nb, tmp_path):
    # the family has one summation only: a file it wrote resumes bit-identically whatever nbody_params.summation says,
    # and a file a stream-K kernel wrote re-pins that kernel (csrc/nbody_records.inl)
    n = 16384
    x0, v0 = nb.ic_plummer(n, seed=8)
    path, path_sk = str(tmp_path / "row.nbody"), str(tmp_path / "sk.nbody")
    with nb.Simulation(n, eps2=1e-4, dt=1e-3) as sim:
        sim.upload(x0, v0).step(3).save(path).step(2)
        straight = sim.download()
    for summation in (nb.NBODY_SUM_TWO_LEVEL, nb.NBODY_SUM_SINGLE):
        with nb.Simulation(n, eps2=1e-4, dt=1e-3, summation=summation) as sim:
            sim.load(path).step(2)
            assert sim.info()["variant_name"].startswith("R_T7_H16") and sim.steps_done() == 5
            for p, q in zip(straight, sim.download()):
                assert np.array_equal(p, q)
    with nb.Simulation(n, eps2=1e-4, dt=1e-3, variant=110, summation=nb.NBODY_SUM_SINGLE) as sim:
        sim.upload(x0, v0).step(3).save(path_sk).step(2)
        straight_sk = sim.download()
    with nb.Simulation(n, eps2=1e-4, dt=1e-3, summation=nb.NBODY_SUM_SINGLE) as sim:     # AUTO = row-split; the file says stream-K
        sim.load(path_sk).step(2)
        assert sim.info()["variant_name"].startswith("T8_NT256_UJ128")
        for p, q in zip(straight_sk, sim.download()):
            assert np.array_equal(p, q)
    with nb.Simulation(n, eps2=1e-4, dt=1e-3) as sim, pytest.raises(nb.NBodyError, match="summation"):
        sim.load(path_sk)                                        # two-level context, single-level file
