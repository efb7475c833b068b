"""Analytic pins for what the reference's fixtures do not cover (last row):
hex8 / tet4 / tri3 element matrices -- symmetry, rigid-body null space, constant-strain
patch test, sum(w*detJ) = volume, and the unit-square quad4 closed form of tmp.txt."""
import numpy as np
import pytest

from fem_b200 import mesh as fm

REF_SHAPES = {
    fm.TRI: np.array([[0, 0], [1, 0], [0, 1]], float),
    fm.QUA: np.array([[-1, -1], [1, -1], [1, 1], [-1, 1]], float),
    fm.TET: np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], float),
    fm.HEX: np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1],
                      [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], float),
}


def distorted(t, seed):
    rng = np.random.default_rng(seed)
    x = REF_SHAPES[t].copy()
    A = np.eye(x.shape[1]) + 0.2 * rng.standard_normal((x.shape[1],) * 2)
    return x @ A.T + 0.15 * rng.uniform(-1, 1, x.shape) + rng.uniform(-3, 3, x.shape[1])


def volume(t, x):
    if t == fm.TRI:
        return 0.5 * np.linalg.det(np.c_[x[1] - x[0], x[2] - x[0]])
    if t == fm.TET:
        return np.linalg.det(np.c_[x[1] - x[0], x[2] - x[0], x[3] - x[0]]) / 6.0
    return None


@pytest.mark.parametrize("t", [fm.TRI, fm.QUA, fm.TET, fm.HEX])
def test_ke_symmetry_rigid_modes_patch(orc, t):
    E, nu = 3.0e10, 0.25
    dim = fm.DIM[t]
    for seed in range(5):
        x = distorted(t, seed)
        K = orc.form_el_ke(t, x, E, nu)
        scale = np.abs(K).max()
        assert np.allclose(K, K.T, atol=1e-13 * scale)
        # rigid translations and infinitesimal rotations are in the null space
        modes = [np.tile(np.eye(dim)[d], len(x)) for d in range(dim)]
        if dim == 2:
            modes.append(np.c_[-x[:, 1], x[:, 0]].reshape(-1))
        else:
            for a, b in ((0, 1), (1, 2), (2, 0)):
                r = np.zeros_like(x); r[:, a] = -x[:, b]; r[:, b] = x[:, a]
                modes.append(r.reshape(-1))
        for mvec in modes:
            assert np.abs(K @ mvec).max() <= 1e-11 * scale * max(1.0, np.abs(mvec).max())
        # constant strain field: K u = integral B^T sigma, and stress at every ip = D*eps
        G = 1e-3 * np.random.default_rng(seed + 10).standard_normal((dim, dim))
        u = x @ G.T
        eps = 0.5 * (G + G.T)
        if dim == 2:
            voigt = np.array([eps[0, 0], eps[1, 1], 2 * eps[0, 1]])
        else:
            voigt = np.array([eps[0, 0], eps[1, 1], eps[2, 2], 2 * eps[0, 1], 2 * eps[1, 2], 2 * eps[2, 0]])
        cp = len(voigt)
        D = np.zeros((cp, cp))
        orc.lib().orc_dmat(dim, orc.C.c_double(E), orc.C.c_double(nu), D.ctypes.data_as(orc._dp))
        sig = orc.calc_el_stress(t, x, u.reshape(-1), E, nu)
        assert np.allclose(sig, D @ voigt, rtol=1e-10, atol=1e-10 * np.abs(D @ voigt).max())
        # strain energy: u^T K u = V * eps:D:eps for a constant strain state
        pts, w = orc.sampts(t)
        vol = 0.0
        for ip in range(len(w)):
            dN = orc.shape_d(t, pts[ip])
            vol += w[ip] * np.linalg.det(dN @ x)
        if volume(t, x) is not None:
            assert np.isclose(vol, volume(t, x), rtol=1e-12)
        assert np.isclose(u.reshape(-1) @ K @ u.reshape(-1), vol * voigt @ D @ voigt, rtol=1e-10)
        # nodal extrapolation reproduces a constant field
        sn = orc.recover_nodal_stress(t, sig)
        assert np.allclose(sn, np.tile(D @ voigt, (len(x), 1)), rtol=1e-10, atol=1e-10 * np.abs(D @ voigt).max())


def test_unit_square_closed_form(orc):
    """tmp.txt rows 6-7: E=3640 nu=.3 unit square -> 2100, 875, -1400, -1050, 350, 175."""
    x = np.array([[0, 2], [0, 1], [1, 1], [1, 2]], float)   # element 1 of cohesive_test_2.inp
    K = orc.form_el_ke(fm.QUA, x, 3640.0, 0.3)
    assert np.allclose(K[6], [-1400, 175, -1050, -875, 350, -175, 2100, 875], rtol=1e-12)
    assert np.allclose(np.diag(K), 2100.0, rtol=1e-12)


def test_quadrature_tables(orc):
    g = np.sqrt(1.0 / 3.0)
    pts, w = orc.sampts(fm.QUA)
    assert np.array_equal(pts, [[-g, -g], [-g, g], [g, -g], [g, g]]) and np.all(w == 1)
    pts, w = orc.sampts(fm.HEX)
    assert np.array_equal(np.sign(pts), np.sign(REF_SHAPES[fm.HEX])) and np.all(w == 1)
    assert orc.sampts(fm.TET)[1][0] == 1.0 / 6.0 and orc.sampts(fm.TRI)[1][0] == 0.5
    assert orc.sampts(fm.COH)[0][:, 0].tolist() == [-0.5773502691896260, 0.5773502691896260]
    # hex ip k sits next to node k, so N2 is diagonally dominant and N2inv exists
    inv = orc.nodal_stress_inv(fm.HEX)
    N2 = np.array([orc.shape(fm.HEX, j) for j in range(8)])
    assert np.allclose(inv @ N2, np.eye(8), atol=1e-14)
    assert orc.nodal_stress_inv(fm.TRI) is None and orc.nodal_stress_inv(fm.COH) is None


def test_csr_matsetvalues_semantics(orc):
    L = orc.lib()
    h = L.orc_csr_create(orc.C.c_int64(3), 2)
    H = orc.C.c_void_p(h)
    add = lambda r, c, v: L.orc_csr_add(H, orc.C.c_int64(r), orc.C.c_int64(c), orc.C.c_double(v))
    for c in (2, 0, 1, 2, 0):
        add(1, c, 1.5)
    add(0, 2, 0.0)                  # explicit zero creates a nonzero slot
    L.orc_csr_insert(H, orc.C.c_int64(1), orc.C.c_int64(1), orc.C.c_double(7.0))
    rp, ci, va = orc.Csr(h).export(3)
    assert rp.tolist() == [0, 1, 4, 4] and ci.tolist() == [2, 0, 1, 2]
    assert va.tolist() == [0.0, 3.0, 7.0, 3.0]


# ---- loads: EdgeAreaNodes*, ApplyTraction, FormRHS, EnforceBCForce -- unpinned by the reference's
# fixtures (no expected Vec_F anywhere), pinned here analytically
def test_edge_area_nodes_closed_forms(orc):
    hexa = REF_SHAPES[fm.HEX] * [1.0, 2.0, 3.0]          # box 2 x 4 x 6
    want = {1: (12.0, [0, 1, 5, 4]), 2: (24.0, [1, 2, 6, 5]), 3: (12.0, [2, 3, 7, 6]), 4: (24.0, [3, 0, 4, 7]),
            5: (8.0, [0, 1, 2, 3]), 6: (8.0, [4, 5, 6, 7])}
    for side, (a, nodes) in want.items():
        area, slots = orc.edge_area_nodes(fm.HEX, hexa, side)
        assert abs(area - a) <= 1e-14 * a and slots == nodes
    tet = REF_SHAPES[fm.TET]
    for side, (a, nodes) in {1: (0.5, [0, 1, 3]), 2: (np.sqrt(3) / 2, [1, 2, 3]), 3: (0.5, [0, 2, 3]), 4: (0.5, [0, 1, 2])}.items():
        area, slots = orc.edge_area_nodes(fm.TET, tet, side)
        assert abs(area - a) <= 1e-15 and slots == nodes
    qua = REF_SHAPES[fm.QUA] * [1.5, 0.5]
    for side, (a, nodes) in {1: (3.0, [0, 1]), 2: (1.0, [1, 2]), 3: (3.0, [2, 3]), 4: (1.0, [3, 0])}.items():
        area, slots = orc.edge_area_nodes(fm.QUA, qua, side)
        assert abs(area - a) <= 1e-15 and slots == nodes
    tri = REF_SHAPES[fm.TRI]
    assert orc.edge_area_nodes(fm.TRI, tri, 2) == (np.sqrt(2.0), [1, 2])
    assert orc.edge_area_nodes(fm.TRI, tri, 3)[1] == [2, 0]


def test_form_rhs_ramps_tractions_and_bc_override(orc):
    from fem_b200 import meshgen as mg
    import dataclasses
    m = mg.quad_rect(2, 2)                               # unit cells, nodes 1..9 row by row
    bcnode = np.array([[1, 0, 1], [3, 0, 0]], np.int32)  # node 1: x fixed; node 3: both
    bcval = np.array([[0.02, 9.0], [0.0, -0.5]])
    m = dataclasses.replace(m, bcnode=bcnode, bcval=bcval)
    om = orc.Mesh(m)
    t, dt, t_init = 2.0, 0.25, 0.5                       # this interval: [0.5, 0.75]
    fnode = [5, 5, 2, 9]
    fval = [[8.0, -4.0, 0.0, 2.0],                       # window covers the interval: 0.25 / 2 of the load
            [1.0, 1.0, 0.7, 1.7],                        # starts inside: (0.75 - 0.7) / 1.0
            [3.0, 3.0, 0.8, 1.0],                        # starts after t_end: nothing
            [6.0, 0.0, 0.0, 0.5]]                        # ends exactly at t_init: interval of length 0 -> adds 0
    telsd = [[1, 2], [2, 4], [0, 1]]                     # right edge of cell 1 = left edge of cell 2 (nodes 2, 5); element 0 skipped
    tval = [[10.0, 0.0, 0.0, 1.0],                       # tractions: t1, t2 are divided by dt (src/m_global.F90:414-415) -> window [0, 4]
            [0.0, 4.0, 0.0, 1.0],
            [99.0, 99.0, 0.0, 1.0]]
    f = orc.form_rhs(om, fnode, fval, telsd, tval, t_init, dt, t).reshape(-1, 2)
    want = np.zeros((9, 2))
    want[4] += np.array([8.0, -4.0]) * 0.25 / 2.0
    want[4] += np.array([1.0, 1.0]) * (0.75 - 0.7) / (1.7 - 0.7)
    ramp = 0.25 / 4.0
    for node in (1, 4):                                  # 0-based nodes 2 and 5: edge length 1, two nodes
        want[node] += np.array([10.0, 0.0]) * ramp * 1.0 / 2 + np.array([0.0, 4.0]) * ramp * 1.0 / 2
    pen = float(np.float32(1e30))
    want[0, 0] = pen * 0.02 * dt / t                     # EnforceBCForce overrides whatever was summed
    want[2] = [pen * 0.0 * dt / t, pen * -0.5 * dt / t]
    assert np.allclose(f, want, rtol=1e-15, atol=0.0)
    assert f[1, 0] == 10.0 * ramp / 2 and f[8, 0] == 0.0


def test_form_rhs_hex_face_traction_total(orc):
    from fem_b200 import meshgen as mg
    m = mg.hex_cube(3, clamp=False)
    om = orc.Mesh(m)
    # traction on side 6 (top face) of every top-layer element: total force = traction * area of the top
    top = [e + 1 for e in range(m.nels) if np.allclose(m.coords[m.elem_nodes(e)[4:] - 1, 2], m.coords[:, 2].max())]
    telsd = [[e, 6] for e in top]
    tval = [[0.0, 0.0, -7.0, 0.0, 1.0]] * len(top)
    f = orc.form_rhs(om, [], [], telsd, tval, 0.0, 1.0, 1.0).reshape(-1, 3)
    area = np.ptp(m.coords[:, 0]) * np.ptp(m.coords[:, 1])
    assert abs(f[:, 2].sum() + 7.0 * area) <= 1e-12 * 7.0 * area and np.abs(f[:, :2]).max() == 0.0
