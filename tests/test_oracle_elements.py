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
