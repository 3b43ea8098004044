#!/usr/bin/env python
"""bench.py -- BASELINE.json metric on BASELINE config C2 (configs[1]): primary-ray closest-hit,
1920x1080, 1 spp, procedural 1 002 528-triangle height-field, user program programs/primary.cl
through RD::TraceRays. One "step" = one TraceRays frame (W*H primary rays).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N > 1 is launched by torchrun (one rank per GPU): the scene is replicated, the frame is split into
contiguous runs of 8x4 screen tiles (one run per rank), and the RGBA8 framebuffer is gathered on
rank 0 with NCCL (the only collective). Scaling is "strong" (the frame is fixed).
Prints ONE JSON line on rank 0. See DESIGN.md "Measurement" for every field's definition.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

WIDTH, HEIGHT, GRID_N = 1920, 1080, 708
WORKLOAD = "C2: primary-ray closest-hit, 1920x1080, 1 spp, height-field n=708 (1002528 triangles), programs/primary.cl"
NODE_BYTES, TRI_BYTES, RAY_BYTES, HIT_BYTES = 80, 48, 32, 16


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU code (oracle/_ref: its bvh.cpp + radiance.cl/pbr.cl, the
    same user program compiled as C++) on the host cores, each step a bounded sample of the C2 frame."""
    if rank != 0:
        return
    pkg, ref = entry.load_package(), entry.load_oracle()
    if not ref.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libradiance_ref.so not built"}))
        return
    scenes = pkg.scenes
    cores = os.cpu_count() or 1
    b = scenes.heightfield_scene(GRID_N, WIDTH, HEIGHT)
    t0 = time.time()
    tl = scenes.build_reference_scene(ref, b)
    build_s = time.time() - t0
    n = WIDTH * HEIGHT
    scratch, image = np.zeros(n * 4, np.float32), np.zeros(n * 4, np.uint8)
    rt = scenes.rtprop(0, 1, 1, 0)
    host = [np.array(rt).reshape(1), scratch, image, np.array(b.camera).reshape(1), np.array(b.scene).reshape(1),
            b.meshInfoData, b.vertexData, b.indexData, b.uvData, b.normalData, b.materials, None, 0, tl.blob]
    rows = 60   # bounded sample: 60 image rows (115 200 primary rays) through the middle of the frame
    first, count = (HEIGHT // 2 - rows // 2) * WIDTH, rows * WIDTH
    for _ in range(args.warmup):
        ref.raygen("primary", host, n, first, count, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ref.raygen("primary", host, n, first, count, cores)
    dt = (time.perf_counter() - t0) / max(1, args.steps)
    v = count / dt / 1e6
    sample = "%d of %d image rows (%d rays) per step, reference BVH build %.1f s outside the timed region" % (rows, HEIGHT, count, build_s)
    print(json.dumps({"impl": "reference", "metric": "Mrays/s (primary closest-hit, 1080p)", "value": v, "unit": "Mrays/s",
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
                      "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": WORKLOAD, "sample": sample},
                      "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": "reference", "sample": sample},
                      "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
