"""Generate scene / environment fixtures by RUNNING the reference's own host code
(oracle/_ref/libflxref_host.so = its scene loaders, SBVH builder and alias-table builder compiled
in place from /root/reference).  Small fixtures are committed under tests/golden/; larger ones go
to the git-ignored scenes/_cache/ (they still travel to the GPU box with the repo snapshot).

    python tools/make_fixtures.py            # needs /root/reference; run in the build container
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from fluctus_b200 import scene as fscene  # noqa: E402

REF = os.environ.get("FLUCTUS_REF", "/root/reference")
GOLD = os.path.join(ROOT, "tests", "golden")
CACHE = os.path.join(ROOT, "scenes", "_cache")


def main(cache_only=False):
    os.makedirs(GOLD, exist_ok=True)
    os.makedirs(CACHE, exist_ok=True)
    if not cache_only:  # the committed fixtures are only rewritten on request (struct padding bytes are not stable)
        # teapot: reference PLY loader + reference SBVH (committed, ~0.3 MB compressed)
        s = pyoracle.ref_load_scene(os.path.join(REF, "assets/teapot.ply"))
        fscene.save_scene(os.path.join(GOLD, "teapot_sbvh.npz"), s)
        print("teapot:", len(s["tris"]), "tris", len(s["nodes"]), "nodes", len(s["indices"]), "indices")
        env = pyoracle.ref_load_envmap(os.path.join(REF, "assets/env_maps/night.hdr"))
        fscene.save_envmap(os.path.join(GOLD, "night_env.npz"), env)
        print("night.hdr:", env["rgb"].shape)
    for name, rel in (("egyptcat", "assets/egyptcat/egyptcat.obj"), ("psor_cube", "assets/psor/psor-cube.obj"),
                      ("gold_rings", "assets/gold_rings/gold_rings_bark.obj")):
        try:
            s = pyoracle.ref_load_scene(os.path.join(REF, rel))
            fscene.save_scene(os.path.join(CACHE, name + "_sbvh.npz"), s)
            print(name, len(s["tris"]), "tris", len(s["materials"]), "materials", len(s["textures"]), "textures",
                  "types 0x%x" % s["material_types"])
        except Exception as e:  # a missing asset must not block the others
            print(name, "skipped:", e)


if __name__ == "__main__":
    main(cache_only="--cache-only" in sys.argv)
