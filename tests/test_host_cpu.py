"""CPU-side checks: generators, localisation, ABI symbols (no compute calls without a GPU)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from fem_b200 import mesh as fm
from fem_b200 import meshgen as mg


def test_generators_positive_jacobians(orc):
    for m in (mg.hex_cube(3), mg.jitter(mg.hex_cube(3)), mg.tet_cube(2), mg.jitter(mg.tet_cube(3)),
              mg.quad_rect(4, 3), mg.cohesive_strip(4, 2)):
        for e in range(m.nels):
            t = int(m.etype[e])
            if t == fm.COH:
                continue
            x = m.coords[m.elem_nodes(e) - 1]
            pts, w = orc.sampts(t)
            for ip in range(len(w)):
                assert np.linalg.det(orc.shape_d(t, pts[ip]) @ x) > 0
    m = mg.hex_cube(4)
    assert (m.nels, m.nnds) == (64, 125)
    assert m.elem_nodes(0).tolist() == [1, 2, 7, 6, 26, 27, 32, 31]
    t = mg.tet_cube(3)
    assert t.nels == 6 * 27
    vol = sum(abs(np.linalg.det(np.c_[(t.coords[t.elem_nodes(e)[1:] - 1] - t.coords[t.elem_nodes(e)[0] - 1]).T])) / 6
              for e in range(t.nels))
    assert np.isclose(vol, 27.0)
    s = mg.cohesive_strip(5, 2)
    assert (s.etype == fm.COH).sum() == 5 and s.nnds == 2 * 6 * 3
    e = s.nels - 5
    x = s.coords[s.elem_nodes(e) - 1]
    assert np.allclose(x[0], x[3]) and np.allclose(x[1], x[2]) and x[1, 0] - x[0, 0] == 1.0


def test_inp_roundtrip(tmp_path, golden):
    m = fm.read_inp(golden["decks"]["two_quads_qs"])
    p = tmp_path / "x.inp"
    fm.write_inp(m, str(p))
    m2 = fm.read_inp(str(p))
    for f in ("etype", "eptr", "eind", "emat", "enlmat", "coords", "mat", "cohprops", "bcnode", "bcval", "fnode", "fval"):
        assert np.array_equal(getattr(m, f), getattr(m2, f)), f
    assert (m.t, m.dt, m.nodal_bw) == (m2.t, m2.dt, m2.nodal_bw)


def test_localize_matches_reference_numbering():
    g = mg.quad_rect(4, 4)
    epart = mg.slab_partition(g, 2)
    tot = 0
    for rank in range(2):
        l = fm.localize(g, epart, rank)
        tot += l.nels
        assert np.all(np.diff(l.nl2g) > 0)                       # ascending global ids
        assert np.array_equal(l.coords, g.coords[l.nl2g - 1])
        ge = np.nonzero(epart == rank)[0]
        for k, e in enumerate(ge):                                 # ascending global element order
            assert np.array_equal(l.nl2g[l.elem_nodes(k) - 1], g.elem_nodes(e))
    assert tot == g.nels
    assert fm.petsc_ownership(10, 3, 0) == (0, 4) and fm.petsc_ownership(10, 3, 2) == (7, 10)


def test_abi_exports_every_declared_symbol():
    from fem_b200 import _lib
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "femx.h")).read()
    declared = set(re.findall(r"\b(femx_[a-z_A-Z0-9]+)\s*\(", hdr))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    L = _lib.lib()                     # loads and binds every symbol
    assert L.femx_version() >= 100


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from fem_b200 import api
    with pytest.raises(api.FemxError, match="no CUDA device"):
        api.Femx()
