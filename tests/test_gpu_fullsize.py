"""BASELINE.json configurations at their FULL sizes, through the C-ABI, checked by size-independent
properties (the oracle cannot finish these in seconds; small-size oracle parity is in
test_gpu_parity.py):
  C2 hex8 100^3   -- two independent algorithms (symmetric pair-block kernel, node-row kernel) agree to
                     1e-12, rigid-body modes give zero internal force, a checksum of K u;
  C3 tet4 94^3x6  -- simplex rows kernel (auto): rigid-body modes, constant stress of a linear field;
  C4 quad4 2x(1000x1000) + 1000 cohesive -- fused rows kernel (auto): K rigid translation of each block, Jacobian = K off the
                     interface, residual finite.
"""
import numpy as np
import pytest

from fem_b200 import api, meshgen as mg

pytestmark = pytest.mark.gpu


def _rel(a, b):
    s = max(np.abs(b).max(), 1e-300)
    return np.abs(a - b).max() / s


def test_c2_hex100_cross_variant_and_rigid_modes():
    m = mg.jitter(mg.hex_cube(100, clamp=False))
    u = mg.displacement_field(m)
    out = {}
    for var in (12, 20, 9):
        with api.Femx(assembly_variant=var) as f:
            f.set_mesh(m).build_plan()
            vals, fint = f.assemble_K_fint(u)
            assert f.nnz == 245438109 and len(vals) == f.nnz
            if var != 9:
                _, f_tr = f.assemble_K_fint(np.tile([1.0, -2.0, 0.5], (m.nlnds, 1)))
                assert np.abs(f_tr).max() <= 1e-10 * np.abs(vals).max()
            out[var] = (vals, fint)
    # three independent algorithms (symmetric pair blocks with FMA, DMMA tiles with in-register y sums, per-row recompute in
    # the general fused rows kernel): different summation trees, same matrix
    scale = np.abs(out[9][0]).max()
    for var in (12, 20):
        assert np.abs(out[var][0] - out[9][0]).max() <= 1e-12 * scale, var
        assert _rel(out[var][1], out[9][1]) <= 1e-12, var
        assert np.isfinite(out[var][0]).all()


def test_c3_tet94_rigid_modes_and_constant_stress():
    m = mg.tet_cube(94)
    assert m.nels == 4983504
    with api.Femx() as f:
        f.set_mesh(m).build_plan()
        x = m.coords
        vals, f_tr = f.assemble_K_fint(np.tile([1.0, -2.0, 0.5], (m.nlnds, 1)))
        assert len(vals) == 113806467
        scale = np.abs(vals).max()
        assert np.abs(f_tr).max() <= 1e-10 * scale
        rot = np.c_[-x[:, 1], x[:, 0], np.zeros(len(x))]
        _, f_rot = f.assemble_K_fint(rot)
        assert np.abs(f_rot).max() <= 1e-9 * scale * np.abs(x).max()
        G = 1e-3 * np.array([[1, 2, 0], [0, -1, 3], [2, 0, 1.0]])
        s = f.stress(x @ G.T)
        assert np.abs(s - s[0]).max() <= 1e-9 * np.abs(s[0]).max()


def test_c4_quad_cohesive_strip_jacobian_residual():
    m = mg.cohesive_strip(1000, 1000)
    assert m.nels == 2001000
    with api.Femx(quirks=api.Q_UEL) as f:
        f.set_mesh(m).build_plan()
        nn = m.nlnds
        vals, f_tr = f.assemble_K_fint(np.tile([0.3, -0.7], (nn, 1)))
        scale = np.abs(vals).max()
        assert np.abs(f_tr).max() <= 1e-10 * scale          # elastic blocks: rigid translation; cohesive Ke = 0
        u = np.zeros((nn, 2))
        u[nn // 2:, 1] = 1e-3                                # open the interface by 1e-3
        J = f.CalcJacobian(vals, u, m.dt)
        assert np.isfinite(J).all()
        changed = np.flatnonzero(J != f.apply_kbc(vals.copy()))
        assert 0 < len(changed) <= 64 * 1000                 # only the 8x8 blocks of the 1000 cohesive elements
        du = np.zeros(2 * nn)
        R = f.CalcResidual(J, u, du, np.zeros(2 * nn), m.dt, m.t)
        assert np.isfinite(R).all() and np.abs(R).max() > 0.0
