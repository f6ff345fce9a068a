"""Experiment (GPU box): how much would the traversal kernels gain if the extension / shadow queues were
ordered for coherence?  Takes the bench workload to a steady-state iteration, reorders the queue on the host by
several keys and times k_ext / k_shadow on each ordering (the kernels only read rays and write hit records, so
re-running them is idempotent)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
import fluctus_b200 as fx  # noqa: E402
from fluctus_b200 import abi  # noqa: E402


def morton3(c, bits):
    out = np.zeros(len(c), np.uint64)
    for b in range(bits):
        for a in range(3):
            out |= ((c[:, a].astype(np.uint64) >> b) & 1) << np.uint64(3 * b + a)
    return out


def keys(o, d, lo, hi):
    q = lambda bits: np.clip(((o - lo) / (hi - lo) * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    octant = ((d[:, 0] < 0).astype(np.uint64) | ((d[:, 1] < 0).astype(np.uint64) << 1) | ((d[:, 2] < 0).astype(np.uint64) << 2))
    return {
        "octant": octant,
        "cell8^3": morton3(q(3), 3),
        "octant+cell8^3": (octant << np.uint64(9)) | morton3(q(3), 3),
        "cell16^3+octant": (morton3(q(4), 4) << np.uint64(3)) | octant,
        "octant+cell16^3": (octant << np.uint64(12)) | morton3(q(4), 4),
        "morton30": morton3(q(10), 10),
        "octant+morton30": (octant << np.uint64(30)) | morton3(q(10), 10),
    }


def main():
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    teapot, scene, env, settings = bench.load_workload()
    tr = fx.Tracer(bench.WIDTH, bench.HEIGHT, num_tasks=bench.NUM_TASKS, device=0, stream=stream.cuda_stream, **settings)
    tr.init(scene, env)
    c = tr.clctx
    c.setPixelIndex(0)
    c.render(0, reset=True)
    c.render(int(os.environ.get("EXP_ITERS", "30")))
    c.enqueueWfLogicKernel(firstIteration=False)
    c.enqueueWfRaygenKernel()
    c.enqueueWfMaterialKernels()
    cnt = c.getCounters()
    st = c.exportState()
    for name, qname, oslot, dslot, launch in (("ext", "extension", "orig", "dir", c.enqueueWfExtRayKernel),
                                              ("shadow", "shadow", "shadowOrig", "shadowDir", c.enqueueWfShadowRayKernel)):
        n = int(cnt[qname + "Queue"])
        q0 = c.readQueue(qname, n)
        o = abi.state_field(st, oslot)[q0].astype(np.float64)
        d = abi.state_field(st, dslot)[q0].astype(np.float64)
        lo, hi = o.min(0), o.max(0) + 1e-6
        orders = {"as produced": np.arange(n), "by path id": np.argsort(q0, kind="stable"),
                  "shuffled": np.random.default_rng(1).permutation(n)}
        for k, v in keys(o, d, lo, hi).items():
            orders[k] = np.argsort(v, kind="stable")
        print("%s queue: %d rays" % (name, n))
        for label, perm in orders.items():
            c.writeQueue(qname, q0[perm])
            ts = []
            for _ in range(4):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); launch(); e1.record(stream); e1.synchronize()
                ts.append(e0.elapsed_time(e1))
            print("  %-18s %.3f ms  (%.0f Mrays/s)" % (label, min(ts[1:]), n / min(ts[1:]) * 1e-3))
        c.writeQueue(qname, q0)


if __name__ == "__main__":
    main()
