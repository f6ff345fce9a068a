"""Bit-identity check for the albedo-gamma table (MaterialsDev::gammaLut, flx_shade.cuh): deterministic microkernel
renders of a textured + constant-albedo scene, dumped as raw bits.  Run once per build and compare:

    make -C fluctus_b200/csrc EXTRA=-DFLX_GAMMA_PER_PATH  (after touching the .cu files) ; python tools/check_gamma_lut.py a.npy
    make -C fluctus_b200/csrc                             (default build: table)         ; python tools/check_gamma_lut.py b.npy
    python tools/check_gamma_lut.py a.npy b.npy     -> "identical" or the number of differing words
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))

if len(sys.argv) == 3:
    a, b = np.load(sys.argv[1]), np.load(sys.argv[2])
    d = int((a.view(np.uint32) != b.view(np.uint32)).sum())
    print("identical (%d words)" % a.size if d == 0 else "DIFFERENT: %d of %d words" % (d, a.size))
    sys.exit(0 if d == 0 else 1)

from conftest import GOLDEN, cached_scene  # noqa: E402
from fluctus_b200 import Tracer, scene as fs  # noqa: E402
from fluctus_b200.context import BUF_PIXELS  # noqa: E402
from fluctus_b200.standin import kitchen_standin  # noqa: E402

teapot = fs.load_scene(os.path.join(GOLDEN, "teapot_sbvh.npz"))
env = fs.load_envmap(os.path.join(GOLDEN, "night_env.npz"))
out = []
for name in ("egyptcat", "kitchen"):
    if name == "egyptcat":
        scene = cached_scene("egyptcat")
        b = scene["bounds"]
        c = 0.5 * (b[0] + b[1])
        view = dict(cam_pos=(float(c[0]), float(c[1]), float(b[1][2] + 1.5 * (b[1][2] - b[0][2]))),
                    light_pos=(float(c[0]) + 1.0, float(b[1][1]) + 1.0, float(c[2])))
    else:
        scene, view = kitchen_standin(teapot, grid=(3, 2, 2))
    W, H = 160, 120
    tr = Tracer(W, H, num_tasks=W * H, max_bounces=6, use_env_map=True, use_area_light=True, **view)
    tr.init(scene, env)
    tr.renderSingle(8)
    out.append(tr.clctx.readPixels(BUF_PIXELS).copy().reshape(-1))
    tr.close()
np.save(sys.argv[1], np.concatenate(out))
print("wrote", sys.argv[1], float(np.concatenate(out).mean()))
