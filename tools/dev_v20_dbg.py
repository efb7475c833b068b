#!/usr/bin/env python
"""Phase-skipping diagnostics of hex_mma_kernel (variant 21 = clock-accounting build): device time with parts switched off."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import api, meshgen as mg
m = mg.jitter(mg.hex_cube(100, clamp=False))
u = mg.displacement_field(m)
names = {0: "all on", 1: "no geometry", 2: "no DMMA", 4: "no z-carry", 8: "no ring stores", 16: "no gather loads", 32: "no value stores", 64: "no f",
         3: "no geometry, no DMMA", 15: "MMA warps idle", 112: "gather warps idle", 127: "everything off"}
for dbg, nm in names.items():
    with api.Femx(assembly_variant=21, dbg=dbg) as f:
        f.set_mesh(m).build_plan()
        d_u = api.DeviceBuffer(f, u.nbytes).upload(u)
        d_v = api.DeviceBuffer(f, 8 * f.nnz)
        d_f = api.DeviceBuffer(f, 8 * f.nrows)
        ts = []
        for _ in range(5):
            f._ck(f.L.femx_assemble_K_fint(f.h, d_u.ptr, d_v.ptr, d_f.ptr, api.DEVICE))
            ts.append(f.L.femx_last_kernel_ms(f.h))
        print(f"dbg {dbg:3d} {nm:24s}: {min(ts[1:]):.4f} ms", file=sys.stderr, flush=True)
