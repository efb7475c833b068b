"""Pins the CPU oracle against the reference's own golden vectors."""
import dataclasses
import math

import numpy as np

from fem_b200 import mesh as fm


def _rows(rowptr, colidx, vals, n):
    return [(colidx[rowptr[i]:rowptr[i + 1]].tolist(), vals[rowptr[i]:rowptr[i + 1]]) for i in range(n)]


def _close6(a, b):
    # MatView prints %g (6 significant digits)
    return abs(a - b) <= 6e-6 * max(abs(b), 1e-300) or (abs(b) < 1e-300 and abs(a) < 1e-9)


def test_jacobian_matview_pattern_and_values(golden, orc):
    """examples/tmp.txt:15-33 -- 16x16 Jacobian of examples/cohesive_test_2.inp."""
    g = golden["jacobian_matview_cohesive_test_2"]
    m = fm.read_inp(golden["decks"]["cohesive_test_2"])
    assert m.nels == 3 and m.nnds == 8 and m.etype.tolist() == [2, 2, 5]
    # D2 (src/m_io.F90:305): rank 1 received garbage (~1e-311) material constants for the
    # second quad; emulate with a zero material so rows 4-5 hold the cohesive block alone.
    m = dataclasses.replace(m, mat=np.vstack([m.mat, np.zeros((1, 5))]), emat=np.array([1, 2, 0], np.int32))
    om = orc.Mesh(m)
    K = orc.assemble_K(om, prealloc=m.pdim * (m.nodal_bw + 1), apply_bc=True)
    J = orc.jacobian(om, K, np.zeros(m.nlnds * 2), m.dt, orc.Q_REFERENCE)
    rows = _rows(*J.export(16), 16)
    # sparsity pattern: bit-exact for all 16 rows
    for i in range(16):
        assert rows[i][0] == [c for c, _ in g[i]], f"row {i} pattern"
    # values: rows whose contributions are not polluted by D2 / the 2-rank MatCopy
    for i in (0, 1, 2, 3, 4, 5, 6, 7):
        for (c, gv), v in zip(g[i], rows[i][1]):
            assert _close6(v, gv), (i, c, v, gv)
    # penalty diagonal is the single-precision literal
    assert rows[0][1][0] == float(np.float32(1.0e30))
    # D4 signature: unsigned 2*e*sigma_max/delta_n on every node pair of the normal column
    assert math.isclose(rows[4][1][0], 2 * math.e * 100.0 / 0.01, rel_tol=1e-14)


def test_jacobian_uel_tangent_is_symmetric(golden, orc):
    m = fm.read_inp(golden["decks"]["cohesive_test_2"])
    rng = np.random.default_rng(3)
    u = 1e-3 * rng.standard_normal(m.nlnds * 2)
    e = 2
    xe = m.coords[m.elem_nodes(e) - 1]
    S = orc.apply_stiff_1(xe, m.elem_nodes(e), u, m.dt, m.cohprops[0], orc.Q_UEL)
    assert np.allclose(S, S.T, rtol=1e-13, atol=1e-9)
    Sref = orc.apply_stiff_1(xe, m.elem_nodes(e), u, m.dt, m.cohprops[0], orc.Q_TANGENT_PRECEDENCE)
    assert not np.allclose(S, Sref)


def test_elastic_golden_vtk_stress(golden, orc):
    """examples/SNES_test/elastic: sigma = (105, 245, 0), u_y = 0.05*y (analytic too)."""
    coords = np.array([[0, 2], [0, 1], [0, 0], [1, 2], [1, 1], [1, 0]], float)
    m = fm.from_uniform(2, fm.QUA, [[1, 2, 5, 4], [2, 3, 6, 5]], coords, E=3640.0, nu=0.3)
    u = np.zeros((6, 2))
    u[:, 1] = 0.05 * coords[:, 1]
    s, cnt = orc.stress(orc.Mesh(m), u.reshape(-1))
    assert cnt.tolist() == [1, 2, 1, 1, 2, 1]
    for rank in ("elastic_vtk_rank0", "elastic_vtk_rank1"):
        gv = golden[rank]
        for p, sg, dg in zip(gv["points"], gv["stress"], gv["disp"]):
            n = int(np.argmin(np.abs(coords - np.array(p[:2])).sum(1)))
            assert np.allclose(s[n], sg, atol=5e-7)          # F0.6 print
            assert np.allclose(u[n], dg[:2], atol=5e-7)
    assert np.allclose(s, [105.0, 245.0, 0.0], rtol=1e-13, atol=1e-11)


def test_uel_traction_known_answer(golden, orc):
    """examples/CZM_for_instability_test.log:70-112: per-ip nodal forces of the UEL element
    (nodes (0,1),(1,1),(1,1),(0,1); props 100, .01, .01, 1, 0).  The opening that produced
    them is back-solved from T_n*w*det."""
    props = np.array([100.0, 0.01, 0.01, 1.0, 0.0, 0.0])
    xe = np.array([[0, 1], [1, 1], [1, 1], [0, 1]], float)
    for blk in (0, 2):
        ip1, ip2 = np.array(golden["uel_ip_forces"][blk]), np.array(golden["uel_ip_forces"][blk + 1])
        N1 = orc.shape(orc.COH, 0)[0]
        target_tn = ip1[1] / (N1 * 0.5)
        # solve e*smax*x*exp(-x) = Tn for x = gap/dn on the rising branch
        x = 0.0
        for _ in range(200):
            fx = math.e * 100.0 * x * math.exp(-x) - target_tn
            x -= fx / (math.e * 100.0 * (1 - x) * math.exp(-x))
        u = np.zeros(8)
        u[5] = u[7] = x * 0.01       # top nodes (3, 4) open in +y
        total, per_ip = orc.apply_tract_1(xe, [1, 2, 3, 4], u, 0.01, props, orc.Q_UEL, per_ip=True)
        for got, want in ((per_ip[0], ip1), (per_ip[1], ip2)):
            assert np.allclose(got[1::2], want[1::2], rtol=2e-13, atol=0)
            assert np.all(np.abs(got[0::2]) < 1e-30) and np.all(np.abs(want[0::2]) < 1e-30)
        assert np.allclose(total, per_ip[0] + per_ip[1], rtol=1e-15)
        # D8: reference keeps only the last ip
        last = orc.apply_tract_1(xe, [1, 2, 3, 4], u, 0.01, props, orc.Q_FORCE_LAST_IP)
        assert np.array_equal(last, per_ip[1])


def test_seplaw_matches_closed_form_and_its_own_tangent(orc):
    """src/m_seplaw.F90 vs the published Xu-Needleman form (Cohesive_element.f:214-229)."""
    p = np.array([100.0, 0.01, 0.02, 0.7, 0.3, 0.0])
    smax, dn, dt_, q, r = p[:5]
    phi = math.e * smax * dn
    rng = np.random.default_rng(0)
    for _ in range(20):
        gn, gt = rng.uniform(-0.002, 0.02), rng.uniform(-0.02, 0.02)
        T = orc.seplaw_1_tract(p, [gn, gt], [0, 0], 0.01)
        ex = math.exp(-gt * gt / dt_ ** 2)
        Tn = phi / dn * math.exp(-gn / dn) * (gn / dn * ex + (1 - q) / (r - 1) * (1 - ex) * (r - gn / dn))
        Tt = 2 * phi / dt_ * (q + (r - q) / (r - 1) * gn / dn) * math.exp(-gn / dn) * ex * gt / dt_
        assert np.allclose(T, [Tn, Tt], rtol=1e-13, atol=1e-12)
        S = orc.seplaw_1_stiff(p, [gn, gt], [0, 0], 0.01)
        h = 1e-7
        for j in range(2):
            d = np.zeros(2); d[j] = h
            fd = (orc.seplaw_1_tract(p, np.array([gn, gt]) + d, [0, 0], .01) -
                  orc.seplaw_1_tract(p, np.array([gn, gt]) - d, [0, 0], .01)) / (2 * h)
            # stiff(1,2)=stiff(2,1) is imposed by the reference (:90-91); only the diagonal and
            # dTn/dgt are the analytic derivative in general.
            assert math.isclose(S[j, j], fd[j], rel_tol=1e-5, abs_tol=1e-3)
    # viscosity terms: zeta*smax*vgap/dn on T_n, zeta*smax/dn/dt on S11
    pz = p.copy(); pz[5] = 0.25
    T0 = orc.seplaw_1_tract(p, [1e-3, 0], [0.5, 0], 0.01)
    T1 = orc.seplaw_1_tract(pz, [1e-3, 0], [0.5, 0], 0.01)
    assert math.isclose(T1[0] - T0[0], 0.25 * 100 * 0.5 / 0.01, rel_tol=1e-12)
    S0 = orc.seplaw_1_stiff(p, [1e-3, 0], [0.5, 0], 0.01)
    S1 = orc.seplaw_1_stiff(pz, [1e-3, 0], [0.5, 0], 0.01)
    assert math.isclose(S1[0, 0] - S0[0, 0], 0.25 * 100 / 0.01 / 0.01, rel_tol=1e-10)


def test_decks_parse(golden):
    m = fm.read_inp(golden["decks"]["lin_two_quads_qs"])
    assert (m.nels, m.nnds, len(m.mat), len(m.bcnode), len(m.fnode)) == (6, 9, 2, 3, 2)
    assert m.etype.tolist() == [2, 1, 1, 1, 1, 2]
    assert m.elem_nodes(0).tolist() == [4, 5, 8, 7] and m.elem_nodes(5).tolist() == [2, 3, 6, 5]
    assert m.bcnode.tolist() == [[1, 0, 0], [7, 0, 0], [2, 1, 1]]
    assert m.fnode.tolist() == [9, 3] and m.fval[0].tolist() == [-10.0e10, 0, 0, 0.1]
    m2 = fm.read_inp(golden["decks"]["two_quads_qs"])
    assert (m2.nels, m2.nnds, len(m2.cohprops)) == (8, 12, 1)
    assert m2.enlmat.tolist() == [0, 0, 0, 0, 0, 0, 1, 1]
    assert m2.cohprops[0].tolist() == [100.0, 0.01, 0.01, 1.0, 0.0, 0.0]
