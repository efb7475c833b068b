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


def test_vtk_writer_reproduces_golden_files_byte_for_byte(golden, orc):
    """WriteOutput (src/m_io.F90:480-555): the reference's own two output files must come back byte-identical from
    the arrays they hold (F0.3 / F0.6 edit descriptors, no leading zero, `-.000000` for the solver's negative
    round-off), and from the oracle's stress of the displacement they hold."""
    for r in (0, 1):
        text = golden[f"elastic_vtk_rank{r}_text"]
        pts, cells, types, stress, disp = fm.read_vtk(text)
        m = fm.from_uniform(2, fm.QUA, [[c + 1 for c in cell] for cell in cells], pts[:, :2], E=3640.0, nu=0.3)
        assert types == [fm.VTK_ID[fm.QUA]] * m.nels
        assert fm.write_vtk(m, stress, disp[:, :2]) == text                      # the writer itself: byte-identical
        s, _ = orc.stress(orc.Mesh(m), disp[:, :2].reshape(-1))
        # recomputed stress: identical text up to the sign of a rounded-to-zero shear (the file's own displacements
        # carry the solver's round-off only as `-.000000`)
        unsign = lambda t: t.replace("-.000000", ".000000")
        assert unsign(fm.write_vtk(m, s, disp[:, :2])) == unsign(text)
    assert [fm.fortran_f0(v, 3) for v in (0.0, -0.0, 0.5, -0.25, 12.3456, -1e-9, 999.9996)] == \
        [".000", "-.000", ".500", "-.250", "12.346", "-.000", "1000.000"]
    # 3-D: no padding column, six stress components
    h = mg.hex_cube(1, clamp=False)
    t = fm.write_vtk(h, np.arange(48.0).reshape(8, 6), np.zeros((8, 3))).split("\n")
    assert t[4] == "POINTS 8 double" and t[13] == "CELLS 1 9" and t[14] == "8 0 1 3 2 4 5 7 6" and t[16] == "12"
    assert t[18] == "SCALARS STRESS FLOAT 6" and t[20] == ".000000 1.000000 2.000000 3.000000 4.000000 5.000000"


def test_loads_roundtrip_and_distribution_sum_to_global_rhs(tmp_path, orc):
    """.inp grammar incl. forces and tractions (ReadDistForces / ReadDistTractions, src/m_io.F90:330-475): a deck
    written and read back is unchanged, and the per-rank FormRHS vectors of a 3-way split add up to the 1-rank one."""
    import dataclasses
    g = mg.jitter(mg.hex_cube(4))
    fnode, fval, telsd, tval = mg.random_loads(g, nfrcs=17, ntrcs=23, seed=3)
    telsd[telsd[:, 0] == 0, 0] = 1
    g = dataclasses.replace(g, fnode=fnode, fval=fval, telsd=telsd, tval=tval)
    p = tmp_path / "loads.inp"
    fm.write_inp(g, str(p))
    back = fm.read_inp(str(p))
    assert np.array_equal(back.fnode, fnode) and np.array_equal(back.fval, fval)
    assert np.array_equal(back.telsd, telsd) and np.array_equal(back.tval, tval)
    f_all = orc.form_rhs(orc.Mesh(dataclasses.replace(g, bcnode=g.bcnode[:0], bcval=g.bcval[:0])), fnode, fval, telsd, tval, 0.0, g.dt, g.t)
    epart = mg.slab_partition(g, 3)
    total = np.zeros_like(f_all)
    seen_f = seen_t = 0
    for r in range(3):
        loc = fm.localize(g, epart, r)
        loc = dataclasses.replace(loc, bcnode=loc.bcnode[:0], bcval=loc.bcval[:0])
        fn, fv, te, tv = fm.distribute_loads(g, epart, 3, r)
        seen_f += len(fn); seen_t += len(te)
        # nodal forces may name nodes this rank does not hold (VecSetValue is global): FormRHS only needs the global dof
        total += orc.form_rhs(orc.Mesh(loc), fn, fv, te, tv, 0.0, g.dt, g.t)
    assert seen_f == len(fnode) and seen_t == len(telsd)
    assert np.allclose(total, f_all, rtol=1e-13, atol=1e-13 * np.abs(f_all).max())


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
