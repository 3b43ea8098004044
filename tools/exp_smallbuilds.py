"""Developer tool: build time of a scene of many small meshes (C4: 4096 patch meshes of 2450 triangles) through the Python
mirror, split into the bottom-level builds and the instance build, for the fused one-launch small build against the
multi-launch one (RD_BUILD_UNFUSED=1) and for several side-stream counts (RD_BUILD_STREAMS).
usage: python tools/exp_smallbuilds.py [side]   (side x side patch meshes, default 64)"""
import os
import sys
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

pkg = entry.load_package()
rd, scenes = pkg.rd, pkg.scenes
side = int(sys.argv[1]) if len(sys.argv) > 1 else 64
bundle = scenes.groups_scene(side=side)
print("meshes", len(bundle.meshes), "triangles", bundle.triangle_count(), flush=True)
rays = scenes.pinhole_rays(bundle.camera, stride=16)


def run(label, env):
    for k in ("RD_BUILD_UNFUSED", "RD_BUILD_STREAMS", "RD_BUILD_PAGEABLE"):
        os.environ.pop(k, None)
    os.environ.update(env)
    plt = rd.Platform(0)   # a fresh platform: the side streams are created on first use
    for rep in range(3):
        plt.sync()
        t0 = time.perf_counter()
        blas = [rd.BuildAccelStruct(plt, rd.Mesh(v, t)) for (v, t) in bundle.meshes]
        t1 = time.perf_counter()
        inst = [rd.Instance(m, sbt, custom, blas[mi]) for (mi, m, sbt, custom) in bundle.instances]
        tlas = rd.BuildAccelStruct(plt, inst)
        plt.sync()
        t2 = time.perf_counter()
        hitA, hitB = rd.trace_host_rays(plt, tlas, rays)
        crc = zlib.crc32(hitA.tobytes() + hitB.tobytes())
        nodes = sum(b.stats()["nodes"] for b in blas[:64])
        print("%-28s rep %d  blas calls %.1f ms  tlas (waits for the builds) %.1f ms  total %.1f ms  nodes[:64] %d  hits crc %08x" % (
            label, rep, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t2 - t0) * 1e3, nodes, crc), flush=True)
        for b in blas:
            b.free()


def run_device(label, env):
    """the same builds from arrays already in HBM (rd_blas_build_device): what the per-mesh uploads cost"""
    for k in ("RD_BUILD_UNFUSED", "RD_BUILD_STREAMS", "RD_BUILD_PAGEABLE"):
        os.environ.pop(k, None)
    os.environ.update(env)
    plt = rd.Platform(0)
    bufs = []
    for (v, t) in bundle.meshes:
        v = np.ascontiguousarray(v, np.float32); t = np.ascontiguousarray(t, np.uint32)
        dv, di = rd.CreateBuffer(plt, v.nbytes), rd.CreateBuffer(plt, t.nbytes)
        rd.WriteBuffer(plt, dv, v.nbytes, v); rd.WriteBuffer(plt, di, t.nbytes, t)
        bufs.append((dv.device_ptr(), v.shape[0], di.device_ptr(), t.shape[0], dv, di))
    L = rd.lib()
    for rep in range(3):
        plt.sync()
        t0 = time.perf_counter()
        blas = []
        for (pv, nv, pi, nt, _a, _b) in bufs:
            out = rd.C.c_void_p()
            rd._check(L.rd_blas_build_device(plt.handle, pv, nv, pi, nt, rd.C.byref(out)))
            blas.append(rd.BottomAccelStruct(plt, out))
        t1 = time.perf_counter()
        nodes = sum(b.stats()["nodes"] for b in blas)
        t2 = time.perf_counter()
        print("%-28s rep %d  blas calls %.1f ms  all resolved after %.1f ms  nodes %d" % (label, rep, (t1 - t0) * 1e3, (t2 - t0) * 1e3, nodes), flush=True)
        for b in blas:
            b.free()


if len(sys.argv) > 2 and sys.argv[2] == "once":   # one cold run under the environment as given (first build of a fresh process)
    env = {k: os.environ[k] for k in ("RD_BUILD_UNFUSED", "RD_BUILD_STREAMS", "RD_BUILD_PAGEABLE") if k in os.environ}
    run("cold: %s prewarm %s" % (env, os.environ.get("RD_POOL_PREWARM_MB", "default")), env)
    sys.exit(0)
for ns in (8, 16, 32, 64):
    run("fused, %d streams" % ns, {"RD_BUILD_STREAMS": str(ns)})
run_device("device arrays, fused, 32", {"RD_BUILD_STREAMS": "32"})
run_device("device arrays, fused, 64", {"RD_BUILD_STREAMS": "64"})
run("fused, 32, pageable copies", {"RD_BUILD_STREAMS": "32", "RD_BUILD_PAGEABLE": "1"})
run("unfused, 8 streams", {"RD_BUILD_UNFUSED": "1", "RD_BUILD_STREAMS": "8"})
