"""femx_stress_noalloc through the C-ABI against femx_stress (collected last on purpose).

The entry point has no counterpart in the reference (include/femx.h says so): it is femx_stress with the per-(element,
node) scratch in a file-scope device array of FEMX_STRESS_NOALLOC_MAX_DOUBLES doubles, for callers that may not allocate.
Checked here: same results as femx_stress on two calls in a row (different displacements, so the second call must not see
the first one's scratch), across contexts that share the array (hex8 -> tet4 -> hex8 again), both output layouts, and the
refusals: a mesh beyond the static size, pageable host pointers.
"""
import ctypes as C

import numpy as np
import pytest

from fem_b200 import _lib, api, meshgen as mg

pytestmark = pytest.mark.gpu

TOL_FIELD = 1e-10          # the stress tolerance of tests/test_gpu_parity.py


def _close(a, b):
    scale = np.abs(b).max()
    return bool(np.all(np.abs(a - b) <= TOL_FIELD * np.maximum(np.abs(b), scale)))


def _noalloc(fx, d_u, d_s, u, layout=api.ROW_MAJOR):
    cp, n = fx.mesh.cpdim, fx.mesh.nlnds
    d_u.upload(np.ascontiguousarray(u, np.float64).reshape(-1))
    d_s.upload(np.full(n * cp, np.nan))                       # a slot the call skips would stay NaN
    fx._ck(fx.L.femx_stress_noalloc(fx.h, d_u.ptr, d_s.ptr, layout))
    out = d_s.download(count=n * cp)
    return out.reshape(cp, -1).T if layout == api.COL_MAJOR else out.reshape(-1, cp)


def test_noalloc_stress_matches_femx_stress_on_consecutive_calls():
    meshes = [mg.jitter(mg.hex_cube(7, clamp=False)), mg.jitter(mg.tet_cube(5)), mg.cohesive_strip(9, 4)]
    ctxs = []
    try:
        for m in meshes:
            fx = api.Femx()
            fx.set_mesh(m).build_plan()
            ctxs.append((fx, fx.device_buffer(8 * m.pdim * m.nlnds), fx.device_buffer(8 * m.cpdim * m.nlnds)))
        launches = 0
        for rep in range(2):                                   # two rounds over the contexts: every call finds another mesh's scratch
            for fx, d_u, d_s in ctxs:
                for call in range(2):                          # two calls in a row on one context
                    u = mg.displacement_field(fx.mesh) * (1.0 + 0.37 * call + 0.11 * rep)
                    want = fx.stress(u)
                    before = fx.launches
                    got = _noalloc(fx, d_u, d_s, u)
                    launches += fx.launches - before
                    assert np.all(np.isfinite(got)), (rep, call)
                    assert _close(got, want), (rep, call)
                    assert np.array_equal(got, want), (rep, call)    # same kernels, same order: the bits agree too
                u = mg.displacement_field(fx.mesh)
                assert np.array_equal(_noalloc(fx, d_u, d_s, u, api.COL_MAJOR), fx.stress(u, layout=api.COL_MAJOR))
        assert launches >= 2 * 2 * 2 * len(ctxs)               # the CUDA kernels ran (>= 2 launches per call)
    finally:
        for fx, d_u, d_s in ctxs:
            d_u.free(); d_s.free(); fx.close()


def test_noalloc_stress_refuses_what_it_cannot_do_without_allocating():
    L = _lib.lib()
    cap = int(L.femx_stress_noalloc_max_doubles())
    assert cap == 2097152
    with api.Femx() as fx:
        m = mg.hex_cube(36, clamp=False)                       # 46656 hex8 * 8 nodes * 6 components > the static array
        assert m.eptr[-1] * m.cpdim > cap
        fx.set_mesh(m).build_plan()
        d_u, d_s = fx.device_buffer(8 * 3 * m.nlnds), fx.device_buffer(8 * 6 * m.nlnds)
        rc = fx.L.femx_stress_noalloc(fx.h, d_u.ptr, d_s.ptr, api.ROW_MAJOR)
        assert rc == 5                                         # FEMX_ERR_UNSUPPORTED
        msg = fx.L.femx_last_error(fx.h).decode()
        assert str(cap) in msg and str(int(m.eptr[-1]) * 6) in msg, msg
        d_u.free(); d_s.free()
    with api.Femx() as fx:
        m = mg.hex_cube(3, clamp=False)
        fx.set_mesh(m).build_plan()
        u, s = np.zeros(3 * m.nlnds), np.zeros(6 * m.nlnds)
        rc = fx.L.femx_stress_noalloc(fx.h, u.ctypes.data_as(C.c_void_p), s.ctypes.data_as(C.c_void_p), api.ROW_MAJOR)
        assert rc != 0 and "device" in fx.L.femx_last_error(fx.h).decode()
        # the context is still usable afterwards
        assert np.all(fx.stress(np.zeros((m.nlnds, 3))) == 0.0)
