#!/usr/bin/env python
"""Wall time of the one-time integer pass (femx_build_plan) on the BASELINE configurations, device pass vs host pass.
usage: plan_bench.py [c2] [c3] [c4] [c5] [host]      (FEMX_PLAN_TIMING=1 prints the stages of the device pass)"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import api, meshgen as mg

CASES = {
    "c2": ("C2 hex8 100^3", lambda: mg.jitter(mg.hex_cube(100, clamp=False))),
    "c3": ("C3 tet4 94^3x6", lambda: mg.jitter(mg.tet_cube(94))),
    "c4": ("C4 quad4 2x1000^2 + 1000 cohesive", lambda: mg.jitter(mg.cohesive_strip(1000, 1000))),
    "c5": ("C5 hex8 200^3", lambda: mg.jitter(mg.hex_cube(200, clamp=False))),
}
want = [a for a in sys.argv[1:] if a in CASES] or ["c2", "c3", "c4"]
for key in want:
    name, gen = CASES[key]
    m = gen()
    for host in ((False, True) if "host" in sys.argv else (False,)):
        with api.Femx(host_plan=host) as f:
            t0 = time.time(); f.set_mesh(m); t1 = time.time()
            best = 1e9
            for _ in range(3):
                f.build_plan()
                best = min(best, f.L.femx_plan_seconds(f.h))
            print(json.dumps({"config": name, "plan": "host (plan.cpp)" if host else "device (plan_gpu.cu)", "elements": int(m.nels), "nnz": int(f.nnz),
                              "set_mesh_s": round(t1 - t0, 4), "build_plan_s": round(best, 4)}), flush=True)
