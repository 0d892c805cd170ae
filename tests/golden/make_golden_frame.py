"""make_golden_frame.py -- regression fixture for the headless frame.

    make -C oracle && python tests/golden/make_golden_frame.py

The frame is produced by oracle_render (oracle/nbody_oracle.c), the CPU restatement of the
reference's shaders (TestProject/simulation_visualization.cpp:15-56), on the reference's own
initial conditions (fill_with_random4, N_BODIES = 8192, constants.h:13).  No OpenGL stack
exists on this box or the GPU box, so there is no frame of the reference itself to pin it
against: PARITY UNPINNED -- this file guards the restatement (and through the GPU test the
CUDA kernels) against regressions only.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle as o  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
x = o.ref_fill_with_random4(8192) if o.ref_lib() is not None else o.ic_reference(8192)
assert np.array_equal(x, o.ic_reference(8192))
img = o.render(x)
np.savez_compressed(os.path.join(OUT, "frame_n8192.npz"), rgba=img, lit_pixels=np.int64((img[:, :, 3] > 0).sum()))
print("frame written:", int((img[:, :, 3] > 0).sum()), "lit pixels")
