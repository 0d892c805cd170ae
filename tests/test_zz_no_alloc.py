"""nbody_step_inplace_noalloc_f4 (include/nbody_b200.h; csrc/nbody_noalloc.inl): the in-place step that allocates
nothing -- its kernel's scratch is a static device array of the library.  No counterpart in the reference.
Checked through the C ABI against the existing entry point
nbody_step_inplace_f4 on two calls in a row, with the tolerances of tests/test_parity_gpu.py.

This file is collected last on purpose: a failure here hides nothing."""
import ctypes as C

import numpy as np
import pytest

TOL_A = 1e-5     # as tests/test_parity_gpu.py: per-body norm-relative
TOL_X = 1e-6
TOL_V = 1e-5
EPS2, DT = 0.002, 0.1   # TestProject/constants.h:25-26


def norm_rel_err(a, b):
    a = np.asarray(a, np.float64)[:, :3]
    b = np.asarray(b, np.float64)[:, :3]
    den = np.linalg.norm(b, axis=1)
    den = np.where(den > 0, den, 1.0)
    return float(np.max(np.linalg.norm(a - b, axis=1) / den))


def initial_bodies(nb, n, seed):
    x = np.zeros((n, 4), np.float32)
    assert nb.lib.nbody_ic_uniform(x.ctypes.data_as(C.c_void_p), n, seed, 1e5, 1e5, 1e9) == 0
    return x


def ptr(t):
    return C.c_void_p(t.data_ptr())


def test_exported_and_refuses_sizes_beyond_the_static_scratch(nb):
    # host-only part (runs on CPU too): the limit is stated and checked before any CUDA call
    limit = nb.lib.nbody_noalloc_max_bodies()
    assert limit == 16384
    assert nb.lib.nbody_step_inplace_noalloc_f4(1, 1, 1, limit + 1, EPS2, DT, None) == -1
    assert b"static scratch" in nb.lib.nbody_last_error()
    assert nb.lib.nbody_step_inplace_noalloc_f4(None, 1, 1, 16, EPS2, DT, None) == -1
    assert nb.lib.nbody_step_inplace_noalloc_f4(1, 1, 1, 0, EPS2, DT, None) == -1
    assert nb.lib.nbody_step_inplace_noalloc_f4(1, 1, 1, 16, 0.0, DT, None) == -1


# 16 384 = the maximum (BASELINE config 2); 10 000: ragged, split tiles; 3 000 / 700: small systems.  The existing
# entry point picks its kernel by size (small-N, row-split or stream-K family); the allocation-free one always runs the
# stream-K kernel, so the two agree to summation rounding, not bit for bit
@pytest.mark.gpu
@pytest.mark.parametrize("n", [16384, 10000, 3000, 700])
def test_two_calls_in_a_row_match_the_existing_entry_point(cuda, nb, n):
    torch = cuda
    x0 = initial_bodies(nb, n, seed=2024 + n)
    xr, vr, ar = torch.from_numpy(x0).cuda(), torch.zeros(n, 4, device="cuda"), torch.zeros(n, 4, device="cuda")
    xn, vn, an = xr.clone(), vr.clone(), ar.clone()
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    free_before = None
    for call in range(2):
        assert nb.lib.nbody_step_inplace_f4(ptr(xr), ptr(ar), ptr(vr), n, EPS2, DT) == 0, nb.lib.nbody_last_error()
        if call == 1:   # the first call may load the module; the second must leave the device's free memory alone
            free_before = torch.cuda.mem_get_info()[0]
        rc = nb.lib.nbody_step_inplace_noalloc_f4(ptr(xn), ptr(an), ptr(vn), n, EPS2, DT, C.c_void_p(stream.cuda_stream))
        assert rc == 0, nb.lib.nbody_last_error()
        stream.synchronize()
        if call == 1:
            assert torch.cuda.mem_get_info()[0] >= free_before
        a_ref, a_new = ar.cpu().numpy(), an.cpu().numpy()
        assert np.isfinite(a_new).all() and np.abs(a_new[:, :3]).max() > 0
        assert norm_rel_err(a_new, a_ref) <= TOL_A, (call, norm_rel_err(a_new, a_ref))
        assert norm_rel_err(vn.cpu().numpy(), vr.cpu().numpy()) <= TOL_V
        assert norm_rel_err(xn.cpu().numpy(), xr.cpu().numpy()) <= TOL_X
        assert torch.equal(xn[:, 3], xr[:, 3]) and float(an[:, 3].abs().max()) == 0.0   # mass untouched, A.w = 0


@pytest.mark.gpu
def test_shared_scratch_is_clean_between_sizes_and_streams(cuda, nb):
    # the array keeps the last call's contents: a large step, then a small one on the default stream, then the
    # large one again must give what a fresh run gives
    torch = cuda
    n_big, n_small = 16384, 1000
    xb0 = torch.from_numpy(initial_bodies(nb, n_big, 7)).cuda()
    xs0 = torch.from_numpy(initial_bodies(nb, n_small, 8)).cuda()

    def run(x0, n, stream):
        x, v, a = x0.clone(), torch.zeros(n, 4, device="cuda"), torch.zeros(n, 4, device="cuda")
        torch.cuda.synchronize()
        assert nb.lib.nbody_step_inplace_noalloc_f4(ptr(x), ptr(a), ptr(v), n, EPS2, DT, stream) == 0, nb.lib.nbody_last_error()
        torch.cuda.synchronize()
        return x, v, a

    first = run(xb0, n_big, None)
    small = run(xs0, n_small, None)
    s = torch.cuda.Stream()
    again = run(xb0, n_big, C.c_void_p(s.cuda_stream))
    for p, q in zip(first, again):
        assert torch.equal(p, q)
    xs, vs, as_ = xs0.clone(), torch.zeros(n_small, 4, device="cuda"), torch.zeros(n_small, 4, device="cuda")
    assert nb.lib.nbody_step_inplace_f4(ptr(xs), ptr(as_), ptr(vs), n_small, EPS2, DT) == 0
    assert norm_rel_err(small[2].cpu().numpy(), as_.cpu().numpy()) <= TOL_A
