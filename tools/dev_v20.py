#!/usr/bin/env python
"""Development check of assembly variant 20 (hex_mma_kernel): oracle parity on the small hex8 cases of
tests/test_gpu_parity.py, then the 100^3 cube against variant 12 with device times."""
import dataclasses, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import api, mesh as fm, meshgen as mg
from oracle import pyoracle as orc
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_gpu_parity import _permuted, _l_shaped, close, TOL_MAT

VAR = int(os.environ.get("FEMX_VAR", "20"))
def small():
    g = mg.jitter(mg.hex_cube(7))
    cases = {
        "cube3": mg.jitter(mg.hex_cube(3, clamp=False)),
        "box_9x5x11": mg.hex_block(9, 5, 11),
        "cube17_chunks": mg.jitter(mg.hex_cube(17, clamp=False)),
        "box_20x30x4": mg.jitter(mg.hex_block(20, 30, 4)),
        "permuted": _permuted(mg.jitter(mg.hex_cube(6, clamp=False)), 3),
        "l_shaped": _l_shaped(8),
        "rank2of3": fm.localize(g, mg.slab_partition(g, 3), 2),
        "two_materials": dataclasses.replace(mg.jitter(mg.hex_cube(5, clamp=False)),
                                             mat=np.array([[3e10, .25, 0, 0, 0], [7e9, .25, 0, 0, 0]]),
                                             emat=(np.arange(125) % 2 + 1).astype(np.int32)),
    }
    ok = True
    for name, m in cases.items():
        m = dataclasses.replace(m, bcnode=np.zeros((0, 4), np.int32), bcval=np.zeros((0, 3)))
        u = mg.displacement_field(m)
        om = orc.Mesh(m)
        rp_o, ci_o, va_o = orc.assemble_K(om).export(om.ndof)
        f_o = orc.fint(om, u.reshape(-1))
        for lz in (0, 25, 3):
            with api.Femx(assembly_variant=VAR, lz=lz) as f:
                f.set_mesh(m).build_plan()
                v, fi = f.assemble_K_fint(u)
                k = f.assemble_K()
            sc = np.abs(va_o).max()
            ev = np.abs(v - va_o).max() / sc
            ef = np.abs(fi - f_o).max() / np.abs(f_o).max()
            good = close(v, va_o, TOL_MAT) and close(fi, f_o, TOL_MAT) and np.array_equal(k, v)
            print(f"{name:16s} lz={lz:2d} errK={ev:.2e} errf={ef:.2e} konly_same={np.array_equal(k, v)} {'OK' if good else 'FAIL'}", flush=True)
            ok = ok and good
    return ok

def big(n=100):
    m = mg.jitter(mg.hex_cube(n, clamp=False))
    u = mg.displacement_field(m)
    out = {}
    for var in ((VAR,) if "only" in sys.argv else (12, VAR)):
        with api.Femx(assembly_variant=var) as f:
            f.set_mesh(m).build_plan()
            d_u = api.DeviceBuffer(f, u.nbytes).upload(u)
            d_v = api.DeviceBuffer(f, 8 * f.nnz)
            d_f = api.DeviceBuffer(f, 8 * f.nrows)
            ts = []
            for _ in range(8):
                f._ck(f.L.femx_assemble_K_fint(f.h, d_u.ptr, d_v.ptr, d_f.ptr, api.DEVICE))
                ts.append(f.L.femx_last_kernel_ms(f.h))
            out[var] = (d_v.download(), d_f.download(), min(ts[2:]))
            print(f"variant {var}: {min(ts[2:]):.4f} ms (all {['%.3f' % t for t in ts]})", flush=True)
    if "only" in sys.argv: return True
    sc = np.abs(out[12][0]).max()
    ev = np.abs(out[VAR][0] - out[12][0]).max() / sc
    ef = np.abs(out[VAR][1] - out[12][1]).max() / np.abs(out[12][1]).max()
    print(f"100^3: variant {VAR} vs 12: errK={ev:.2e} errf={ef:.2e}")
    return ev <= 1e-12 and ef <= 1e-12

if __name__ == "__main__":
    ok = small() if "nosmall" not in sys.argv else True
    if "big" in sys.argv:
        ok = big() and ok
    print("DEV", "OK" if ok else "FAILED")
    sys.exit(0 if ok else 1)
