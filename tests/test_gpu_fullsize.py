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


def _sampled_rows_vs_oracle(orc, m, rowptr, colidx, vals, fint, u, nodes):
    """The oracle assembles just the elements around each sampled node (FormElKE per element, MatSetValues(ADD) in ascending
    element order, src/main.F90:157-166) and the node's pdim matrix rows and internal-force entries are compared with the
    full-size CUDA result at 1e-12 (relative to the largest entry of the matrix / vector)."""
    pd = m.pdim
    nen = (m.eptr[1:] - m.eptr[:-1]).astype(np.int64)
    el_of = np.repeat(np.arange(m.nels, dtype=np.int64), nen)
    order = np.argsort(m.eind, kind="stable")                 # incidence: ascending node, ascending element inside
    nptr = np.searchsorted(m.eind[order], np.arange(1, m.nlnds + 2))
    vscale, fscale = np.abs(vals).max(), np.abs(fint).max()
    worst_v = worst_f = 0.0
    for r in nodes:
        rows = {}
        f = np.zeros(pd)
        for p in order[nptr[r]:nptr[r + 1]]:
            e = el_of[p]
            a = p - m.eptr[e]
            en = m.eind[m.eptr[e]:m.eptr[e + 1]].astype(np.int64) - 1
            t = int(m.etype[e])
            if t == 5:                                        # cohesive: zero block, pattern only
                k = np.zeros((4 * pd, 4 * pd))
            else:
                E, nu = m.mat[m.emat[e] - 1, :2]
                k = orc.form_el_ke(t, m.coords[en], E, nu)
            ue = np.asarray(u)[en].reshape(-1)
            gcols = (pd * (m.nl2g[en].astype(np.int64) - 1)[:, None] + np.arange(pd)).reshape(-1)
            for i in range(pd):
                kr = k[a * pd + i]
                d = rows.setdefault(i, {})
                for c, v in zip(gcols, kr):
                    d[c] = d.get(c, 0.0) + v
                f[i] += kr @ ue
        g0 = pd * (int(m.nl2g[r]) - 1)
        for i in range(pd):
            lo, hi = rowptr[g0 + i], rowptr[g0 + i + 1]
            cols = sorted(rows[i])
            assert np.array_equal(colidx[lo:hi], np.array(cols, np.int32)), ("pattern", r, i)
            ref = np.array([rows[i][c] for c in cols])
            worst_v = max(worst_v, np.abs(vals[lo:hi] - ref).max() / vscale)
            worst_f = max(worst_f, abs(fint[g0 + i] - f[i]) / fscale)
    assert worst_v <= 1e-12 and worst_f <= 1e-12, (worst_v, worst_f)


def _sample_nodes(n_nodes, special, count=400, seed=77):
    h = (np.arange(n_nodes, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(seed)) >> np.uint64(33)
    picked = np.argsort(h)[:count]
    return np.unique(np.concatenate([picked, np.asarray(special, np.int64)]))


def _rel(a, b):
    s = max(np.abs(b).max(), 1e-300)
    return np.abs(a - b).max() / s


def test_c2_hex100_cross_variant_and_rigid_modes(orc):
    m = mg.jitter(mg.hex_cube(100, clamp=False))
    u = mg.displacement_field(m)
    out = {}
    # sampled nodes: hash-chosen ones plus the places where tile, chunk and wave boundaries of the two tile kernels fall
    # (8 x 6 tiles of variant 12, 7 x 13 tiles of variant 20, z-chunks of 17 / 51 layers), the mesh faces, edges and corners
    n1 = 101
    idx = lambda i, j, k: (k * n1 + j) * n1 + i
    special = [idx(i, j, k) for i in (0, 6, 7, 8, 15, 16, 48, 49, 55, 56, 98, 99, 100) for j in (0, 5, 6, 12, 13, 14, 25, 26, 77, 78, 90, 91, 100)
               for k in (0, 1, 16, 17, 18, 33, 34, 50, 51, 52, 84, 85, 99, 100)]
    nodes = _sample_nodes(m.nlnds, special, count=300)
    rowptr = colidx = None
    for var in (12, 20, 9):
        with api.Femx(assembly_variant=var) as f:
            f.set_mesh(m).build_plan()
            vals, fint = f.assemble_K_fint(u)
            assert f.nnz == 245438109 and len(vals) == f.nnz
            if var != 9:
                if rowptr is None:
                    rowptr, colidx = f.pattern()
                _sampled_rows_vs_oracle(orc, m, rowptr, colidx, vals, fint, u, nodes)
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


def test_c3_tet94_rigid_modes_and_constant_stress(orc):
    m = mg.tet_cube(94)
    assert m.nels == 4983504
    with api.Femx() as f:
        f.set_mesh(m).build_plan()
        x = m.coords
        u = mg.displacement_field(m)
        vals, fint = f.assemble_K_fint(u)
        rowptr, colidx = f.pattern()
        n1 = 95
        special = [(k * n1 + j) * n1 + i for i in (0, 1, 47, 94) for j in (0, 46, 94) for k in (0, 1, 93, 94)]
        _sampled_rows_vs_oracle(orc, m, rowptr, colidx, vals, fint, u, _sample_nodes(m.nlnds, special, count=250))
        del rowptr, colidx
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


def test_c4_quad_cohesive_strip_jacobian_residual(orc):
    m = mg.cohesive_strip(1000, 1000)
    assert m.nels == 2001000
    with api.Femx(quirks=api.Q_UEL) as f:
        f.set_mesh(m).build_plan()
        nn = m.nlnds
        u = mg.displacement_field(m)
        vals, fint = f.assemble_K_fint(u)
        rowptr, colidx = f.pattern()
        half = nn // 2                                        # interface nodes: last row of the lower block, first of the upper one
        special = [0, 1, 1000, 1001, half - 1001, half - 1000, half - 1, half, half + 1, half + 1000, nn - 1]
        _sampled_rows_vs_oracle(orc, m, rowptr, colidx, vals, fint, u, _sample_nodes(nn, special, count=400))
        del rowptr, colidx
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
