#!/usr/bin/env python
"""Development timing of the simplex rows kernel (variant 11) on C3 (tet4 94^3 x 6, jittered) and a tri3 mesh:
device time per assembly for lanes-per-row G in --g, and the largest relative difference to the two-pass path (variant 1).
FEMX_LIB selects a development build of the library."""
import argparse, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import api, meshgen as mg  # noqa: E402


def run(m, variant, g, reps):
    u = mg.displacement_field(m)
    with api.Femx(assembly_variant=variant, tile=g) as fx:
        fx.set_mesh(m).build_plan()
        d_u = fx.device_buffer(u.nbytes).upload(u)
        d_vals = fx.device_buffer(8 * fx.nnz)
        d_f = fx.device_buffer(8 * fx.nrows)
        ms = []
        for _ in range(reps):
            fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, d_vals.ptr, d_f.ptr, api.DEVICE))
            ms.append(fx.last_kernel_ms)
        return float(np.median(ms[2:])), float(min(ms)), d_vals.download(np.float64, fx.nnz), d_f.download(np.float64, fx.nrows)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tet", type=int, default=94)
    ap.add_argument("--tri", type=int, default=0)
    ap.add_argument("--g", type=int, nargs="*", default=[0])
    ap.add_argument("--reps", type=int, default=9)
    ap.add_argument("--check", type=int, default=1)
    a = ap.parse_args()
    tag = os.path.basename(os.environ.get("FEMX_LIB", "libfemx.so"))
    for name, m in (("tet", mg.jitter(mg.tet_cube(a.tet)) if a.tet else None), ("tri", mg.jitter(mg.tri_rect(a.tri, a.tri)) if a.tri else None)):
        if m is None:
            continue
        ref = run(m, 1, 0, 3) if a.check else None
        for g in a.g:
            med, best, v, f = run(m, 11, g, a.reps)
            d = {"lib": tag, "mesh": name, "nels": int(m.nels), "g": g, "ms_median": med, "ms_min": best}
            if ref:
                d["rel_vals"] = float(np.abs(v - ref[2]).max() / np.abs(ref[2]).max())
                d["rel_f"] = float(np.abs(f - ref[3]).max() / np.abs(ref[3]).max())
            print(json.dumps(d), flush=True)


if __name__ == "__main__":
    main()
