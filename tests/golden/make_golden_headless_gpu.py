"""make_golden_headless_gpu.py -- pins the headless-directory flavour of the step on the reference's OWN code.

That flavour (TestProject/Sim-Without-OpenGL-Integration/kernel.cu:38-125: host float4 bodies, float3 V/A, double
EPS2 / DT) exists in the reference as GPU code only, so its pin is the reference kernel's output on a B200:
oracle/_ref/libref_headless_sm100.so is that file rebuilt for sm_100 by oracle/Makefile (one MSVC-only token fixed
in a /tmp copy).  Run ON A GPU BOX (the .so travels there; /root/reference is not needed):

    python tests/golden/make_golden_headless_gpu.py gpurun_out/headless_ref_gpu.npz

and commit the result as tests/golden/headless_ref_gpu.npz.  Sizes are whole 1024-thread blocks (the launch the
reference makes on this device): N = 1024 is ONE block -- no cross-block race on the positions, summation in
index order, deterministic; N = 4096 is four co-resident blocks.  The partial-tile behaviour of that kernel at other
N (stale shared-memory entries are summed, late threads skip the last tile) is a defect, not a specification, and is
not recorded.  The reduction flavours of the main directory (kernel.cu:163-318 + :406-469; float4 and float3 V/A) are
recorded the same way through oracle/_ref/libref_reduction_sm100.so (the reference's kernels behind a GL-free launch
sequence, oracle/ref_reduction_glue.cu); they have a device synchronisation between summing and integrating, so
they are race-free at any N that is a multiple of 32.  Also prints how far the oracle's restatements and this repo's
engine are from the recorded output.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import oracle as o  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "libref_headless_sm100.so")
REF_RED = os.path.join(ROOT, "oracle", "_ref", "libref_reduction_sm100.so")
STEPS = 5


def ref_simulate():
    lib = C.CDLL(REF)
    f = getattr(lib, "_Z8simulateP6float4P6float3S2_i")   # void simulate(float4*, float3*, float3*, int)
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    f.restype = None
    return lambda x, a, v, n: f(x.ctypes.data, a.ctypes.data, v.ctypes.data, n)


def ref_reduction(stride):
    """the reference's reduction-flavour kernels (kernel.cu:163-318, :406-469) behind oracle/ref_reduction_glue.cu"""
    lib = C.CDLL(REF_RED)
    f = lib.ref_embparallel_step if stride == 4 else lib.ref_embparallel_float3_step
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    f.restype = C.c_int

    def step(x, a, v, n):
        rc = f(x.ctypes.data, v.ctypes.data, a.ctypes.data, n)
        assert rc == 0, f"reference reduction step failed: CUDA error {rc}"
    return step


def run_red(step, n, x0, stride):
    x, v, a = x0.copy(), np.zeros((n, stride), np.float32), np.zeros((n, stride), np.float32)
    step(x, a, v, n)
    first = (x.copy(), v.copy(), a.copy())
    for _ in range(STEPS - 1):
        step(x, a, v, n)
    return first + (x, v, a)


def reduction_flavours(res, report):
    names = ("x1", "v1", "a1", "xk", "vk", "ak")
    for stride in (4, 3):
        ref = ref_reduction(stride)
        for n in (1024, 2048):
            x0 = o.ic_reference(n)                              # TestProject/constants.h: MAX_W 1e9, DT 0.1f
            got = run_red(ref, n, x0, stride)
            assert np.abs(got[2]).max() > 0 and np.isfinite(got[3]).all(), "the reference kernels did not run"
            rep = {"deterministic": bool(all(np.array_equal(p, q) for p, q in zip(got, run_red(ref, n, x0, stride))))}
            res.update({f"red{stride}_{k}_n{n}": g for k, g in zip(names, got) if k != "ak"})
            orc = run_red(lambda x, a, v, m: o.step_mixed(x, a, v), n, x0, stride)
            rep.update({f"oracle_{k}": o.norm_rel_err(p, q) for k, p, q in zip(names, orc, got)})
            try:
                import torch
                import nbody_b200 as nb

                def eng(x, a, v, m):
                    dx, da, dv = (torch.from_numpy(t.reshape(-1)).cuda() for t in (x, a, v))
                    (nb.simulateHeadless_embParallel if stride == 4 else nb.simulateHeadless_embParallel_float3)(dx, da, dv, None, m)
                    x[:], a[:], v[:] = dx.cpu().numpy().reshape(x.shape), da.cpu().numpy().reshape(a.shape), dv.cpu().numpy().reshape(v.shape)
                e = run_red(eng, n, x0, stride)
                rep.update({f"engine_{k}": o.norm_rel_err(p, q) for k, p, q in zip(names, e, got)})
            except Exception as ex:  # noqa: BLE001
                rep["engine_error"] = repr(ex)
            report[f"red{stride}_n{n}"] = rep


def slim(res):
    """shrink what gets committed without losing a bit: drop the float3 flavour's numbers when they equal the float4
    flavour's xyz; drop the initial positions (oracle.ic_reference regenerates them); store positions as their
    difference from the initial ones (exact in fp32 for values this close, mostly zeros, checked here)"""
    out = dict(res)
    for k in [k for k in out if k.startswith("red3_")]:
        twin = out["red4_" + k[5:]]
        assert np.array_equal(out[k], twin[:, :3] if out[k].shape[1] == 3 else twin), k
        del out[k]
    out["red3_equals_red4_xyz"] = np.int32(1)
    for k in [k for k in out if "_n" in k and k.split("_n")[0].split("_")[-1] in ("x1", "xk")]:
        n, headless = int(k.split("_n")[1]), not k.startswith("red4_")
        base = o.ic_reference(n, max_w=1e8) if headless else o.ic_reference(n)
        d = out.pop(k) - base
        assert np.array_equal(base + d, res[k]), k
        out["d_" + k] = d
    for k in [k for k in out if k.startswith("x0_n")]:
        assert np.array_equal(out[k], o.ic_reference(int(k[4:]), max_w=1e8))
        del out[k]
    return out


def positions(g, key, base):
    """inverse of slim() for a position array: initial positions + stored difference (bit-exact)"""
    return base + g["d_" + key]


def run(step, n, x0):
    x, v, a = x0.copy(), np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
    step(x, a, v, n)
    x1, v1 = x.copy(), v.copy()
    for _ in range(STEPS - 1):
        step(x, a, v, n)
    return x1, v1, x, v


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "headless_ref_gpu.npz")
    sim = ref_simulate()
    res, report = {"steps": np.int32(STEPS)}, {}
    for n in (1024, 4096):
        x0 = o.ic_reference(n, max_w=1e8)                       # headless constants.h: MAX_W 1e8
        x1, v1, xk, vk = run(sim, n, x0)
        assert np.abs(v1).max() > 0 and np.isfinite(xk).all(), "the reference kernel did not run"
        again = run(sim, n, x0)
        rep = {"deterministic": bool(all(np.array_equal(p, q) for p, q in zip((x1, v1, xk, vk), again)))}
        res.update({f"x0_n{n}": x0, f"x1_n{n}": x1, f"v1_n{n}": v1, f"xk_n{n}": xk, f"vk_n{n}": vk})
        # the oracle's restatement against it
        X, V3 = x0.copy(), np.zeros((n, 3), np.float32)
        o.step_headless_f3(X, V3, 0.002, 0.01)
        rep["oracle_v1"], rep["oracle_x1"] = o.norm_rel_err(V3, v1), o.norm_rel_err(X, x1)
        rep["oracle_x1_bit_equal"] = bool(np.array_equal(X, x1))
        for _ in range(STEPS - 1):
            o.step_headless_f3(X, V3, 0.002, 0.01)
        rep["oracle_vk"], rep["oracle_xk"] = o.norm_rel_err(V3, vk), o.norm_rel_err(X, xk)
        # this repo's engine against it (informational; the test lives in tests/test_parity_gpu.py)
        try:
            import nbody_b200 as nb
            e1 = run(lambda x, a, v, m: nb.simulate_host_f3(x, a, v, m), n, x0)
            rep["engine_v1"], rep["engine_x1"] = o.norm_rel_err(e1[1], v1), o.norm_rel_err(e1[0], x1)
            rep["engine_vk"], rep["engine_xk"] = o.norm_rel_err(e1[3], vk), o.norm_rel_err(e1[2], xk)
        except Exception as e:  # noqa: BLE001
            rep["engine_error"] = repr(e)
        report[f"n{n}"] = rep
    reduction_flavours(res, report)
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    np.savez_compressed(out_path, **slim(res))
    print(json.dumps(report))


if __name__ == "__main__":
    main()
