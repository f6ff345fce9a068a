"""Analysis aid (uses the oracle): share of inner-node visits that falls on the first K inner nodes in
breadth-first order, for extension rays of the benchmark scene.  Decides how many nodes are worth staging in
shared memory.   python tools/top_level_share.py"""
import ctypes as C
import os
import sys
from collections import deque

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from fluctus_b200 import abi, scene as fs  # noqa: E402
import bench  # noqa: E402


def bfs_rank(nodes):
    rank = np.full(len(nodes), -1, np.int64)
    q, k = deque([0]), 0
    leaf = nodes["nPrims"] != 0
    while q:
        i = q.popleft()
        if leaf[i]:
            continue
        rank[i] = k
        k += 1
        q.append(i + 1)
        q.append(int(nodes["iStartOrRight"][i]))
    return rank, k


def main():
    teapot, scene, env, settings = bench.load_workload()
    N, W, H = 1 << 15, 480, 270
    p = abi.make_render_params(W, H, n_tris=len(scene["tris"]), world_radius=fs.world_radius(scene), **settings)
    wf = pyoracle.OracleWavefront(N, W, H)
    wf.set_scene(scene); wf.set_envmap(env); wf.set_params(p.tobytes())
    wf.render(12, True)  # steady state
    L = pyoracle.lib()
    hist = np.zeros(len(scene["nodes"]), np.uint32)
    pyoracle.set_threads(1)
    L.orc_set_node_histogram(hist.ctypes.data_as(C.c_void_p))
    wf.logic(False); wf.raygen(); wf.materials(); wf.ext()
    L.orc_set_node_histogram(None)
    rank, n_inner = bfs_rank(scene["nodes"])
    by_rank = np.zeros(n_inner, np.int64)
    m = rank >= 0
    by_rank[rank[m]] = hist[m]
    total = by_rank.sum()
    print("inner nodes %d, visits %d (%.1f per ray)" % (n_inner, total, total / max(int(wf.counters()['extensionQueue']), 1)))
    cum = np.cumsum(by_rank)
    for k in (63, 127, 255, 447, 511, 895, 1023, 1791, 2047, 3583, 4095, 8191, 16383):
        print("first %5d nodes (%.0f KB): %.1f %% of visits" % (k, k * 64 / 1024, 100.0 * cum[min(k, n_inner) - 1] / total))


if __name__ == "__main__":
    main()
