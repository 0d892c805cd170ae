"""CPU tests of the rows SURVEY.md section 8(f) marks "next": the checker's restatement of the
reference's shaders (f1) and of the reduction flavours' integrator (f2/f4), the exported headless
launch forms, and the argument validation / file handling of the new C-ABI entry points (f3).
No compute call is made on the product here (no GPU)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def body(x, y, z, m):
    return np.array([[x, y, z, m]], np.float32)


# ---- f1: oracle_render = simulation_visualization.cpp:15-56 + fixed-function point rasterisation ----
def test_render_centre_body_lands_on_the_centre_pixel_with_mixed_colour(oracle):
    img = oracle.render(body(0, 0, 0, 1e5))
    ys, xs = np.nonzero(img[:, :, 3])
    # ndc (0,0) -> window (400, 300) -> pixel column 400, GL row 300 = image row 600-1-300
    assert list(zip(ys, xs)) == [(299, 400)]
    assert tuple(img[299, 400]) == (0, 255, 0, 255)          # weight 0 -> vec3(0,1,0) (the shader calls it "blue")
    heavy = oracle.render(body(0, 0, 0, 1e9))
    assert tuple(heavy[299, 400]) == (255, 0, 0, 255)        # weight 1 -> red
    mid = oracle.render(body(0, 0, 0, 1e5 + 0.25 * (1e9 - 1e5)))
    assert tuple(mid[299, 400]) == (64, 191, 0, 255)         # weight 0.25 -> round(63.75), round(191.25)


def test_render_view_volume_is_clipped_and_the_far_plane_fails_gl_less(oracle):
    v = 2e5                                                   # MAX_VIEWX/Y/Z, constants.h:21-23
    for p in ((1.01 * v, 0, 0), (0, -1.01 * v, 0), (0, 0, 1.01 * v), (0, 0, -1.01 * v)):
        assert oracle.render(body(*p, 5e8))[:, :, 3].sum() == 0
    assert oracle.render(body(0, 0, v, 5e8))[:, :, 3].sum() == 0        # depth 1.0 is not < the cleared 1.0
    assert oracle.render(body(0, 0, v, 5e8), depth_test=False)[:, :, 3].sum() == 255
    corner = oracle.render(body(-v, -v, 0, 5e8))
    assert corner[599, 0, 3] == 255                          # ndc (-1,-1) = bottom-left = last image row
    assert oracle.render(body(v, v, 0, 5e8))[:, :, 3].sum() == 0         # ndc +1 -> window 800: outside the viewport


def test_render_depth_test_nearest_wins_and_first_drawn_wins_ties(oracle):
    near, far = [0, 0, -1e5, 1e5], [0, 0, 1e5, 1e9]
    for order in ((near, far), (far, near)):
        img = oracle.render(np.array(order, np.float32))
        assert tuple(img[299, 400]) == (0, 255, 0, 255)      # the nearer (green) body, whatever the draw order
    tie = np.array([[0, 0, 0, 1e5], [0, 0, 0, 1e9]], np.float32)
    assert tuple(oracle.render(tie)[299, 400]) == (0, 255, 0, 255)       # GL_LESS: the second equal depth fails
    assert tuple(oracle.render(tie, depth_test=False)[299, 400]) == (255, 0, 0, 255)   # test off: last drawn wins
    assert tuple(oracle.render(np.array([far, near], np.float32), depth_test=False)[299, 400]) == (0, 255, 0, 255)


def test_render_golden_frame_of_the_reference_initial_conditions(oracle, golden):
    g = golden["frame_n8192"]
    img = oracle.render(oracle.ic_reference(8192))            # fill_with_random4 at N_BODIES (pinned in test_oracle_cpu.py)
    assert img.shape == (600, 800, 4)
    assert np.array_equal(img, g["rgba"])
    assert int((img[:, :, 3] > 0).sum()) == int(g["lit_pixels"])


def test_render_other_frame_sizes(oracle):
    x = oracle.ic_reference(4096)
    small = oracle.render(x, width=64, height=48)
    assert small.shape == (48, 64, 4) and 0 < (small[:, :, 3] > 0).sum() <= 4096
    # bodies fill the inner half of the view volume (MAX_X = MAX_VIEWX / 2): nothing in the outer quarter frame
    assert small[:12].sum() == 0 and small[-12:].sum() == 0 and small[:, :16].sum() == 0 and small[:, -16:].sum() == 0


# ---- f2 / f4: reduction-flavour integrator (kernel.cu:455-463) ---------------------------------------
def test_step_mixed_differs_from_the_float_integrator_only_by_rounding(oracle):
    x0 = oracle.ic_reference(512)
    X1, V1, A1 = x0.copy(), np.zeros_like(x0), np.zeros_like(x0)
    oracle.step_snapshot(X1, A1, V1)
    X2, V2, A2 = x0.copy(), np.zeros_like(x0), np.zeros_like(x0)
    oracle.step_mixed(X2, A2, V2)
    assert np.array_equal(A1, A2)                             # same pair sums
    assert oracle.norm_rel_err(V2, V1) < 2e-7 and oracle.norm_rel_err(X2, X1) < 2e-7
    X3, V3, A3 = x0.copy(), np.zeros((512, 3), np.float32), np.zeros((512, 3), np.float32)
    oracle.step_mixed(X3, A3, V3)                             # float3 layout: same numbers, 12-byte stride
    assert np.array_equal(V3, V2[:, :3]) and np.array_equal(A3, A2[:, :3]) and np.array_equal(X3, X2)


# ---- boundary: exported headless launch forms and option structs -----------------------------------------
def test_headless_forms_of_the_reduction_launchers_are_exported(nb):
    out = subprocess.run(["nm", "-DC", "--defined-only", nb.LIB_PATH], capture_output=True, text=True, check=True).stdout
    for sig in ("simulateHeadless(float4*, float4*, float4*, int)",
                "simulateHeadless_embParallel(float4*, float4*, float4*, float4*, int)",
                "simulateHeadless_embParallel_float3(float4*, float3*, float3*, float3*, int)",
                "simulateHeadless_embParallel_fadl(float4*, float4*, float4*, float4*, int)",
                "simulateHeadless_embParallel_fadl4(float4*, float4*, float4*, float4*, int)"):
        assert sig in out, sig


def test_option_defaults_are_the_reference_constants(nb):
    o = nb.NBodyStepOpts()
    assert nb.lib.nbody_step_opts_default(C.byref(o)) == 0
    assert o.eps2 == np.float32(0.002) and o.dt == np.float32(0.1) and (o.integrator, o.va_float3, o.variant) == (0, 0, 0)
    v = nb.default_view()
    assert (v.width, v.height) == (800, 600)                                   # simulation_visualization.h:8-9
    assert (v.max_x, v.max_y, v.max_z) == (200000.0, 200000.0, 200000.0)       # constants.h:21-23
    assert (v.min_w, v.max_w, v.depth_test) == (100000.0, 1000000000.0, 1)     # constants.h:18-19


def test_new_entry_points_validate_their_arguments(nb, tmp_path):
    o = nb.NBodyStepOpts()
    nb.lib.nbody_step_opts_default(C.byref(o))
    assert nb.lib.nbody_step_inplace_opts(None, None, None, 16, C.byref(o)) == -1
    assert nb.lib.nbody_step_inplace_opts(1, 1, 1, 0, C.byref(o)) == -1
    o.integrator = nb.NBODY_INTEG_F64
    assert nb.lib.nbody_step_inplace_opts(1, 1, 1, 16, C.byref(o)) == -1
    v = nb.default_view()
    assert nb.lib.nbody_render_device(None, 16, C.byref(v), None) == -1
    v.width = 0
    assert nb.lib.nbody_render_device(1, 16, C.byref(v), 1) == -1
    v = nb.default_view(max_w=1e5)                                             # empty weight range
    assert nb.lib.nbody_render_device(1, 16, C.byref(v), 1) == -1
    bad = tmp_path / "not_a_snapshot.bin"
    bad.write_bytes(b"x" * 100)
    with pytest.raises(nb.NBodyError):
        nb.snapshot_info(str(bad))
    with pytest.raises(nb.NBodyError):
        nb.snapshot_info(str(tmp_path / "missing.bin"))


def test_write_ppm_is_host_only_and_drops_alpha(nb, tmp_path):
    img = np.zeros((3, 5, 4), np.uint8)
    img[1, 2] = (10, 20, 30, 255)
    path = tmp_path / "f.ppm"
    nb.write_ppm(str(path), img)
    raw = path.read_bytes()
    assert raw.startswith(b"P6\n5 3\n255\n")
    px = np.frombuffer(raw[len(b"P6\n5 3\n255\n"):], np.uint8).reshape(3, 5, 3)
    assert tuple(px[1, 2]) == (10, 20, 30) and px.sum() == 60
