#!/usr/bin/env python
"""Device-timed lines for the BASELINE.json configurations that are parity-test cases rather than the bench
headline: C3 tet4 94^3x6 (assembly + stress recovery), C4 quad4 2x(1000x1000) + 1000 cohesive (K, Jacobian,
residual).  Prints one JSON object per phase: ms per call (median of --reps), elements/s and algorithmic GB/s
(SURVEY.md 8(d) bytes per element).  Run on a GPU box:  python tools/config_bench.py > gpurun_out/configs.json"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import api, meshgen as mg  # noqa: E402


def timed(fx, call, reps):
    ms = []
    for _ in range(reps):
        call()
        ms.append(fx.last_kernel_ms)
    return float(np.median(ms)), float(min(ms))


def line(cfg, phase, nels, ms, best, bytes_per_el, extra=None):
    d = {"config": cfg, "phase": phase, "elements": int(nels), "ms_median": ms, "ms_min": best,
         "elements_per_s": nels / (ms * 1e-3), "algorithmic_bytes_per_element": bytes_per_el,
         "algorithmic_GBps": nels * bytes_per_el / (ms * 1e-3) / 1e9}
    if extra:
        d.update(extra)
    print(json.dumps(d), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=7)
    ap.add_argument("--tet", type=int, default=94)
    ap.add_argument("--quad", type=int, default=1000)
    ap.add_argument("--g", type=int, default=0, help="simplex rows kernel: lanes per node row (tuning knob, 0 = default)")
    ap.add_argument("--dbg", type=int, default=0)
    args = ap.parse_args()

    # ---- C3: tet4 assembly + stress (two-pass path: element rows -> packed scratch -> ordered gather)
    m = mg.tet_cube(args.tet)
    u = mg.displacement_field(m)
    with api.Femx(tile=args.g, dbg=args.dbg) as fx:
        t0 = time.time()
        fx.set_mesh(m).build_plan()
        plan_s = time.time() - t0
        d_u = fx.device_buffer(u.nbytes).upload(u)
        d_vals = fx.device_buffer(8 * fx.nnz)
        d_f = fx.device_buffer(8 * fx.nrows)
        d_s = fx.device_buffer(8 * 6 * m.nlnds)
        ms, best = timed(fx, lambda: fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, d_vals.ptr, d_f.ptr, api.DEVICE)), args.reps)
        line(f"C3 tet4 {args.tet}^3x6", "K + f_int (auto = variant 11, simplex rows kernel)", m.nels, ms, best, 215.0,
             {"nnz": int(fx.nnz), "plan_build_s": round(plan_s, 2)})
        ms, best = timed(fx, lambda: fx._ck(fx.L.femx_stress(fx.h, d_u.ptr, d_s.ptr, api.ROW_MAJOR, api.DEVICE)), args.reps)
        line(f"C3 tet4 {args.tet}^3x6", "nodal-averaged stress", m.nels, ms, best, 36.5)

    # ---- C4: quad4 + Xu-Needleman strip, one Newton step = Jacobian + residual
    m = mg.cohesive_strip(args.quad, args.quad)
    nn = m.nlnds
    u = np.zeros((nn, 2)); u[nn // 2:, 1] = 1e-3
    with api.Femx(quirks=api.Q_UEL) as fx:
        t0 = time.time()
        fx.set_mesh(m).build_plan()
        plan_s = time.time() - t0
        d_u = fx.device_buffer(u.nbytes).upload(u)
        d_k = fx.device_buffer(8 * fx.nnz)
        d_j = fx.device_buffer(8 * fx.nnz)
        d_f = fx.device_buffer(8 * fx.nrows)
        d_du = fx.device_buffer(8 * fx.nrows).upload(np.zeros(fx.nrows))
        d_fe = fx.device_buffer(8 * fx.nrows).upload(np.zeros(fx.nrows))
        d_r = fx.device_buffer(8 * fx.nrows)
        ms, best = timed(fx, lambda: fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, d_k.ptr, d_f.ptr, api.DEVICE)), args.reps)
        line(f"C4 quad4 2x{args.quad}^2 + {args.quad} coh", "K + f_int (auto = variant 9, fused rows kernel)", m.nels, ms, best, 356.0,
             {"nnz": int(fx.nnz), "plan_build_s": round(plan_s, 2)})
        ms, best = timed(fx, lambda: fx._ck(fx.L.femx_jacobian(fx.h, d_k.ptr, d_u.ptr, float(m.dt), d_j.ptr, api.DEVICE)), args.reps)
        line(f"C4 quad4 2x{args.quad}^2 + {args.quad} coh", "CalcJacobian (copy K, cohesive overlay, BC diagonal)", m.nels, ms, best,
             16.0 * fx.nnz / m.nels)
        ms, best = timed(fx, lambda: fx._ck(fx.L.femx_jacobian_update(fx.h, d_k.ptr, d_u.ptr, float(m.dt), d_j.ptr, api.DEVICE)), args.reps)
        line(f"C4 quad4 2x{args.quad}^2 + {args.quad} coh", "CalcJacobian, Newton iterations after the first (femx_jacobian_update: cohesive blocks + BC diagonal only)",
             m.nels, ms, best, 0.0)
        ms, best = timed(fx, lambda: fx._ck(fx.L.femx_residual(fx.h, d_j.ptr, d_u.ptr, d_du.ptr, d_fe.ptr, float(m.dt), float(m.t),
                                                                 d_r.ptr, api.DEVICE)), args.reps)
        line(f"C4 quad4 2x{args.quad}^2 + {args.quad} coh", "CalcResidual (cohesive forces, BC rows, J du - f)", m.nels, ms, best,
             (12.0 * fx.nnz + 24.0 * fx.nrows) / m.nels)


if __name__ == "__main__":
    main()
