"""GPU parity tests of the rows SURVEY.md section 8(f) marks "next", all through the C ABI:
f1 headless frame (bit-exact bytes vs oracle_render), f2 headless forms of the reduction
launchers, f3 snapshot / resume, f4 float3 velocity / acceleration layout."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_A, TOL_X, TOL_V = 1e-5, 1e-6, 1e-5   # per-body norm-relative, as tests/test_parity_gpu.py


# ---- f1: frame --------------------------------------------------------------------------------------
def gpu_frame(torch, nb, x, view=None):
    view = view or nb.default_view()
    dx = torch.from_numpy(x).cuda()
    out = torch.full((view.height, view.width, 4), 77, dtype=torch.uint8, device="cuda")   # poison: every pixel must be written
    nb.render_device(dx, len(x), out, view)
    return out.cpu().numpy()


def test_frame_matches_golden_and_oracle_bit_for_bit(cuda, nb, oracle, golden):
    x = oracle.ic_reference(8192)
    img = gpu_frame(cuda, nb, x)
    assert np.array_equal(img, golden["frame_n8192"]["rgba"])
    assert np.array_equal(img, oracle.render(x))


@pytest.mark.parametrize("n", [1, 2, 1000, 131072, 1 << 20])
@pytest.mark.parametrize("depth_test", [1, 0])
def test_frame_sizes_and_depth_modes(cuda, nb, oracle, n, depth_test):
    # 1M bodies on 480k pixels: every pixel is contended, the depth / draw-order rule decides all of them
    x = nb.ic_uniform(n, seed=5, max_xyz=2.2e5)            # some bodies outside the view volume (clipped)
    x[: min(n, 64), 3] = np.linspace(0, 2e9, min(n, 64), dtype=np.float32)   # weights outside [0,1]: colour clamps
    view = nb.default_view(depth_test=depth_test)
    img = gpu_frame(cuda, nb, x, view)
    ref = oracle.render(x, depth_test=bool(depth_test))
    assert np.array_equal(img, ref)
    if n >= 1000:
        assert (img[:, :, 3] > 0).sum() > 0


def test_frame_with_duplicate_positions_and_other_window(cuda, nb, oracle):
    x = nb.ic_uniform(5000, seed=9)
    x = np.concatenate([x, x[::-1].copy()])                # every position twice, different draw order
    x[5000:, 3] = x[5000:, 3][::-1]                        # and different masses on the duplicates
    for w, h in ((64, 48), (1920, 1080), (3, 1)):
        for dt_ in (1, 0):
            view = nb.default_view(width=w, height=h, depth_test=dt_, max_z=1.5e5)
            assert np.array_equal(gpu_frame(cuda, nb, x, view),
                                  oracle.render(x, width=w, height=h, max_xyz=(2e5, 2e5, 1.5e5), depth_test=bool(dt_)))


def test_context_frame_follows_the_simulation_and_writes_a_ppm(cuda, nb, oracle, tmp_path):
    x0 = oracle.ic_reference(8192)
    with nb.Simulation(8192) as sim:
        sim.upload(x0)
        f0 = sim.render()
        sim.step(50)
        x, _, _ = sim.download()
        f1 = sim.render()
    assert np.array_equal(f0, oracle.render(x0))
    assert np.array_equal(f1, oracle.render(x))
    assert not np.array_equal(f0, f1)                      # bodies moved
    path = tmp_path / "frame.ppm"
    nb.write_ppm(str(path), f1)
    raw = path.read_bytes()
    assert raw.startswith(b"P6\n800 600\n255\n") and len(raw) == 15 + 800 * 600 * 3


# ---- f2 / f4: headless forms of the reduction launchers, float3 layout ----------------------------------
def test_embparallel_float4_matches_the_reduction_flavour_oracle(cuda, nb, oracle):
    torch = cuda
    n = 3000
    x0 = nb.ic_uniform(n, seed=21)
    X, V, A = x0.copy(), np.zeros_like(x0), np.zeros_like(x0)
    dx, dv, da = torch.from_numpy(x0).cuda(), torch.zeros(n, 4, device="cuda"), torch.full((n, 4), 9.0, device="cuda")
    for _ in range(3):
        oracle.step_mixed(X, A, V)
        nb.simulateHeadless_embParallel(dx, da, dv, None, n)
    a, v, x = da.cpu().numpy(), dv.cpu().numpy(), dx.cpu().numpy()
    assert oracle.norm_rel_err(a, A) <= 3 * TOL_A and oracle.norm_rel_err(v, V) <= 3 * TOL_V and oracle.norm_rel_err(x, X) <= 3 * TOL_X
    assert np.all(a[:, 3] == 0) and np.array_equal(x[:, 3], x0[:, 3])        # kernel.cu:409,433-435


@pytest.mark.parametrize("stride", [4, 3])
@pytest.mark.parametrize("n", [1024, 2048])
def test_embparallel_against_the_reference_kernels_output(cuda, nb, oracle, golden, n, stride):
    # golden = the reference's kernel_reduction[_float3] + inter_block_reduction[_float3] run on a B200
    # (tests/golden/make_golden_headless_gpu.py)
    torch = cuda
    g, k = golden["headless_ref_gpu"], f"red4_%s_n{n}"    # the float3 twin's xyz bits are the same (stored once)
    x0 = oracle.ic_reference(n)
    ref = lambda key: g[k % key][:, :stride] if key[0] != "x" else x0 + g["d_" + k % key]   # noqa: E731
    dx = torch.from_numpy(x0).cuda()
    dv, da = torch.zeros(n * stride, device="cuda"), torch.zeros(n * stride, device="cuda")
    step = nb.simulateHeadless_embParallel if stride == 4 else nb.simulateHeadless_embParallel_float3
    step(dx, da, dv, None, n)
    got = lambda t: t.cpu().numpy().reshape(n, -1)   # noqa: E731
    assert oracle.norm_rel_err(got(da), ref("a1")) <= TOL_A and oracle.norm_rel_err(got(dv), ref("v1")) <= TOL_V
    assert oracle.norm_rel_err(got(dx), ref("x1")) <= TOL_X
    for _ in range(int(g["steps"]) - 1):
        step(dx, da, dv, None, n)
    assert oracle.norm_rel_err(got(dv), ref("vk")) <= TOL_V and oracle.norm_rel_err(got(dx), ref("xk")) <= TOL_X


def test_embparallel_float3_layout(cuda, nb, oracle):
    torch = cuda
    for n in (1, 31, 2049):
        x0 = nb.ic_uniform(n, seed=22)
        v0 = np.random.default_rng(3).uniform(-10, 10, (n, 3)).astype(np.float32)
        X, V, A = x0.copy(), v0.copy(), np.zeros((n, 3), np.float32)
        oracle.step_mixed(X, A, V)
        # guard floats after the arrays: a float4-strided store would trample them
        dv = torch.cat([torch.from_numpy(v0).reshape(-1), torch.full((64,), 123.0)]).cuda()
        da = torch.full((3 * n + 64,), 123.0, device="cuda")
        dx = torch.from_numpy(x0).cuda()
        nb.simulateHeadless_embParallel_float3(dx, da, dv, None, n)
        a, v = da.cpu().numpy(), dv.cpu().numpy()
        assert np.all(a[3 * n:] == 123.0) and np.all(v[3 * n:] == 123.0)
        assert oracle.norm_rel_err(a[:3 * n].reshape(n, 3), A) <= TOL_A
        assert oracle.norm_rel_err(v[:3 * n].reshape(n, 3), V) <= TOL_V
        assert oracle.norm_rel_err(dx.cpu().numpy(), X) <= TOL_X


def test_float3_and_float4_layouts_give_identical_numbers(cuda, nb):
    torch = cuda
    n = 4097
    x0 = nb.ic_uniform(n, seed=23)
    dx4, dv4, da4 = torch.from_numpy(x0).cuda(), torch.zeros(n, 4, device="cuda"), torch.zeros(n, 4, device="cuda")
    dx3, dv3, da3 = torch.from_numpy(x0).cuda(), torch.zeros(n, 3, device="cuda"), torch.zeros(n, 3, device="cuda")
    for _ in range(4):
        nb.simulate_opts(dx4, da4, dv4, n, integrator=nb.NBODY_INTEG_MIXED)
        nb.simulate_opts(dx3, da3, dv3, n, integrator=nb.NBODY_INTEG_MIXED, va_float3=True)
    assert torch.equal(dx4, dx3) and torch.equal(dv4[:, :3], dv3) and torch.equal(da4[:, :3], da3)


def test_opts_step_with_cluster_variants_agrees_with_default(cuda, nb, oracle):
    torch = cuda
    n = 20000
    x0 = nb.ic_uniform(n, seed=24)
    ref = None
    for variant in (0, 60, 64):              # auto | C2 multicast | C2 j-split with DSMEM reduction
        dx, dv, da = torch.from_numpy(x0).cuda(), torch.zeros(n, 4, device="cuda"), torch.zeros(n, 4, device="cuda")
        nb.simulate_opts(dx, da, dv, n, variant=variant)
        a = da.cpu().numpy()
        if ref is None:
            ref = a
        assert oracle.norm_rel_err(a, ref) <= TOL_A, variant   # different summation orders (4e-6 measured)


# ---- f3: snapshot / resume ----------------------------------------------------------------------------------
def test_snapshot_resume_is_bit_identical(cuda, nb, tmp_path):
    n = 5000
    x0, v0 = nb.ic_plummer(n, seed=4)
    path = str(tmp_path / "state.nbody")
    with nb.Simulation(n, dt=0.01) as sim:
        sim.upload(x0, v0).step(10)
        straight = sim.download()
    with nb.Simulation(n, dt=0.01) as sim:
        sim.upload(x0, v0).step(4).save(path)
        assert sim.steps_done() == 4
    assert nb.snapshot_info(path) == {"n_bodies": n, "steps_done": 4, "eps2": np.float32(0.002), "dt": np.float32(0.01)}
    with nb.Simulation(n, dt=0.01) as sim:
        sim.load(path)
        assert sim.steps_done() == 4
        sim.step(6)
        assert sim.steps_done() == 10
        resumed = sim.download()
    for a, b in zip(straight, resumed):
        assert np.array_equal(a, b)
    with nb.Simulation(n + 1) as sim, pytest.raises(nb.NBodyError):
        sim.load(path)                                     # wrong body count


def test_sharded_contexts_resume_from_one_snapshot(cuda, nb, tmp_path):
    n = 3000
    x0 = nb.ic_uniform(n, seed=6)
    path = str(tmp_path / "s.nbody")
    with nb.Simulation(n) as sim:
        sim.upload(x0).step(2).save(path)
        sim.step(1)
        x_ref, v_ref, _ = sim.download()
    shards = [nb.Simulation(n, rank=r, world=3) for r in range(3)]
    try:
        for s in shards:
            s.load(path)
        for s in shards:                                   # all sources are current after load: one local + remote step
            s.shard_local_async()
            s.shard_remote_async()
        xs = np.concatenate([s.download()[0] for s in shards])
        vs = np.concatenate([s.download()[1] for s in shards])
    finally:
        for s in shards:
            s.close()
    o_err = np.abs(xs - x_ref).max() / np.abs(x_ref).max()
    assert o_err < 1e-6 and np.allclose(vs, v_ref, rtol=1e-4, atol=1e-3 * np.abs(v_ref).max())
    with nb.Simulation(n, rank=0, world=2) as s, pytest.raises(nb.NBodyError):
        s.save(path)                                       # a shard holds only its own bodies


# ---- fused exchange: epilogue stores into every rank's next buffer ---------------------------------------------
@pytest.mark.parametrize("world", [2, 3, 8])
def test_fused_exchange_on_one_gpu_matches_unsharded(cuda, nb, oracle, world):
    # G ranks emulated as G contexts of one process on one device, stepped one after the other (no kernel
    # waits on another): each fused step writes its new positions into ALL contexts' next buffers
    torch = cuda
    n = 10001
    x0 = nb.ic_uniform(n, seed=3)
    with nb.Simulation(n) as ref:
        ref.upload(x0).step(3)
        xr, vr, ar = ref.download()
    sims = [nb.Simulation(n, rank=r, world=world) for r in range(world)]
    try:
        with pytest.raises(nb.NBodyError):
            sims[0].shard_step_fused_async()                   # peers not attached yet
        for s in sims:
            s.upload(x0)
        for r, s in enumerate(sims):
            for q, peer in enumerate(sims):
                if q != r:
                    s.shard_attach_ptr(q, *peer.shard_buffers())
        with pytest.raises(nb.NBodyError):
            sims[0].shard_attach_ptr(1, *sims[1].shard_buffers())   # twice
        launches = nb.launch_count()
        for _ in range(3):
            for s in sims:
                s.shard_step_fused_async()
            torch.cuda.synchronize()                           # the cross-rank barrier
        assert nb.launch_count() - launches == 3 * world       # ONE kernel per rank and step
        parts = [s.download() for s in sims]
    finally:
        for s in sims:
            s.close()
    x = np.concatenate([p[0] for p in parts]); v = np.concatenate([p[1] for p in parts]); a = np.concatenate([p[2] for p in parts])
    # same kernel, same source order per target tile only when the stream-K cut coincides: compare to rounding
    # (the sharded-vs-one-GPU bounds of SURVEY 8c: 1e-5 on a and v, 1e-6 on x; three steps, measured 2e-6 .. 3.1e-6)
    assert oracle.norm_rel_err(a, ar) <= 1e-5 and oracle.norm_rel_err(x, xr) <= 1e-6 and oracle.norm_rel_err(v, vr) <= 1e-5


def test_in_kernel_barrier_between_two_gpus(cuda, nb, oracle):
    # the cross-GPU form of the barrier: each rank on its OWN device (kernels that wait for each other must not
    # share a GPU), all steps enqueued back to back with no host synchronisation and no collective in between
    torch = cuda
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (tools/verify_sharded.py --barrier device covers it under torchrun)")
    # moving bodies (Plummer sphere with velocities): a rank that read a peer's slot one step early or late
    # would be off by |v| dt / r ~ 1e-2 in the accelerations, far above the rounding-level tolerance below
    n, steps = 40001, 6
    x0, v0 = nb.ic_plummer(n, seed=5)
    with nb.Simulation(n, dt=0.02) as ref:
        ref.upload(x0, v0).step(steps)
        xr, vr, ar = ref.download()
    sims = [nb.Simulation(n, dt=0.02, rank=r, world=2, device=r) for r in range(2)]
    try:
        for s in sims:
            s.upload(x0, v0)
        sims[0].shard_attach_ptr(1, *sims[1].shard_buffers())
        sims[1].shard_attach_ptr(0, *sims[0].shard_buffers())
        for s in sims:
            s.shard_set_device_barrier(True)
        for _ in range(steps):
            for s in sims:
                s.shard_step_fused_async()
        parts = [s.download() for s in sims]
        assert [s.shard_barrier_timeouts() for s in sims] == [0, 0]
        with pytest.raises(nb.NBodyError):
            sims[0].shard_set_device_barrier(False)            # not once fused steps have run with it
    finally:
        for s in sims:
            s.close()
    x = np.concatenate([p[0] for p in parts]); v = np.concatenate([p[1] for p in parts]); a = np.concatenate([p[2] for p in parts])
    errs = (oracle.norm_rel_err(x, xr), oracle.norm_rel_err(v, vr), oracle.norm_rel_err(a, ar))
    assert errs[0] <= 1e-5 and errs[1] <= 1e-4 and errs[2] <= 1e-4, errs


# ---- launch-bound step loops: CUDA-graph replay gives the same bits as step-by-step launches ---------------
@pytest.mark.parametrize("n", [777, 5000])
def test_graph_replayed_steps_equal_single_steps(cuda, nb, n):
    x0, v0 = nb.ic_plummer(n, seed=8)
    with nb.Simulation(n, dt=0.01) as sim:
        sim.upload(x0, v0)
        for _ in range(71):
            sim.step(1)                                        # 71 separate launches
        one_by_one = sim.download()
    with nb.Simulation(n, dt=0.01) as sim:
        sim.upload(x0, v0).step(3)                             # leaves buffer parity 1: replay must realign
        before = nb.launch_count()
        sim.step(68)                                           # 1 single + 2 x 32 replayed + 3 single
        assert nb.launch_count() - before == 68 and sim.steps_done() == 71
        replayed = sim.download()
    for a, b in zip(one_by_one, replayed):
        assert np.array_equal(a, b)


def test_large_contexts_pick_their_kernel_by_measurement(cuda, nb):
    with nb.Simulation(70000) as sim:                          # >= 64 k bodies: resolved through the tuner
        name = sim.info()["variant_name"]
    rep = nb.autotune_report()
    assert name.startswith("T6_NT256_UJ512")                   # 512-body units from 64 k sources on
    cands = dict(item.split(":") for item in rep.split())
    assert len(cands) >= 4 and all(0.1 < float(ms[:-2]) < 50 for ms in cands.values()), rep
    chosen = [k for k in cands if nb.variants()[int(k)].split(":")[0] == name]
    assert len(chosen) == 1
    # the report rounds to 1 us: the chosen kernel is the fastest OF ITS FAMILY (same tile shape and unit size, hence
    # bit-identical results whichever wins) up to that rounding
    family = [float(ms[:-2]) for k, ms in cands.items() if nb.variants()[int(k)].startswith("T6_NT256_UJ512")]
    assert len(family) == 3 and float(cands[chosen[0]][:-2]) <= min(family) * 1.005 + 0.0015, rep


def test_mid_size_contexts_pick_tile_shape_by_unit_balance(cuda, nb, oracle):
    # 22 000 bodies: 11 tiles of 2048 targets x 128-body units = 1892 units = 12.8 per CTA (13 run) beats
    # 15 tiles of 1536 (the last 2/3 empty) x 256-body units = 1290 units = 8.7 per CTA (9 run);
    # 16 384 = 147 whole tiles of 112 targets: the row-split family, where no tile is shared between CTAs
    with nb.Simulation(22000) as sim:
        assert sim.info()["variant_name"].startswith("T8_NT256_UJ128")
    with nb.Simulation(16384) as sim:
        assert sim.info()["variant_name"].startswith("R_T7_H16")
    with nb.Simulation(50000) as sim:
        assert sim.info()["variant_name"].startswith("T6_NT256_UJ256")
        x0 = nb.ic_uniform(50000, seed=2)
        sim.upload(x0).step(1)
        _, _, a = sim.download()
    idx = np.arange(0, 50000, 977, dtype=np.int32)
    a64 = oracle.accel_f64(x0, idx)
    assert oracle.norm_rel_err(a[idx], a64) <= max(TOL_A, 2 * oracle.norm_rel_err(oracle.accel_f32(x0, idx), a64))
