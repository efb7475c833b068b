#!/usr/bin/env python
"""Small end-to-end pass over every assembly variant, the device plan, stress and the Jacobian on small meshes: written for
compute-sanitizer (closed on this pool), kept as a quick smoke run:  python tools/sanitize_run.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import api, mesh as fm, meshgen as mg

def run(name, m, variants):
    u = mg.displacement_field(m)
    for av in variants:
        with api.Femx(assembly_variant=av) as f:
            f.set_mesh(m).build_plan()
            f.pattern(); f.slot_map()
            v, fi = f.assemble_K_fint(u)
            k = f.assemble_K()
            s = f.stress(u)
            assert np.isfinite(v).all() and np.isfinite(fi).all() and np.array_equal(k, v), (name, av)
    print(name, "ok", flush=True)

run("hex 9x5x11", mg.jitter(mg.hex_block(9, 5, 11)), (0, 12, 20, 9, 1))
run("hex cube 17 (z-chunks)", mg.jitter(mg.hex_cube(17, clamp=False)), (20, 12))
run("tet 5", mg.jitter(mg.tet_cube(5)), (0, 9, 1))
run("quad 9x7", mg.jitter(mg.quad_rect(9, 7)), (0, 1))
run("tri 9x7", mg.jitter(mg.tri_rect(9, 7)), (0, 9, 1))
m = mg.jitter(mg.cohesive_strip(12, 3))
with api.Femx(quirks=api.Q_UEL) as f:
    f.set_mesh(m).build_plan()
    u = mg.displacement_field(m)
    kv = f.apply_kbc(f.assemble_K())
    J = f.CalcJacobian(kv, u, float(m.dt))
    assert np.isfinite(J).all()
print("cohesive strip ok")
print("SMOKE RUN DONE")
