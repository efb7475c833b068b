"""Parity proper: the CUDA path through the C ABI vs the CPU oracle on the same inputs.

Bar (BASELINE.json north_star): sparsity pattern and index maps bit-exact; element,
global-matrix and residual entries <= 1e-12 relative; stress fields <= 1e-10 relative.
"Relative" is |a-b| <= tol * max(|b|, scale) with scale the largest entry of the object
compared, so structurally cancelling entries do not produce false failures.
"""
import dataclasses

import numpy as np
import pytest

from fem_b200 import api, mesh as fm, meshgen as mg

pytestmark = pytest.mark.gpu

TOL_MAT = 1e-12
TOL_FIELD = 1e-10


def close(a, b, tol, scale=None):
    a, b = np.asarray(a), np.asarray(b)
    scale = np.abs(b).max() if scale is None else scale
    return bool(np.all(np.abs(a - b) <= tol * np.maximum(np.abs(b), scale)))


@pytest.fixture(scope="module")
def fx():
    f = api.Femx(quirks=api.Q_REFERENCE)
    yield f
    f.close()


def _distorted(t, n, seed):
    from tests.test_oracle_elements import REF_SHAPES
    rng = np.random.default_rng(seed)
    x = np.repeat(REF_SHAPES[t][None], n, 0)
    d = x.shape[2]
    A = np.eye(d) + 0.2 * rng.standard_normal((n, d, d))
    return np.einsum("nad,ned->nae", x, A) + 0.15 * rng.uniform(-1, 1, x.shape) + rng.uniform(-3, 3, (n, 1, d))


@pytest.mark.parametrize("t", [fm.TRI, fm.QUA, fm.TET, fm.HEX])
def test_element_ke_and_stress(fx, orc, t):
    n = 257
    x = _distorted(t, n, 5)
    rng = np.random.default_rng(9)
    E = rng.uniform(1e3, 5e10, n)
    nu = rng.uniform(0.05, 0.45, n)
    u = 1e-3 * rng.standard_normal(x.shape)
    K = fx.FormElKE(t, x, E, nu)
    s, sn = fx.CalcElStress(t, x, u, E, nu)
    for e in range(n):
        Ko = orc.form_el_ke(t, x[e], E[e], nu[e])
        assert close(K[e], Ko, TOL_MAT), e
        so = orc.calc_el_stress(t, x[e], u[e].reshape(-1), E[e], nu[e])
        assert close(s[e], so, TOL_FIELD), e
        assert close(sn[e], orc.recover_nodal_stress(t, so), TOL_FIELD), e


@pytest.mark.parametrize("quirks", [api.Q_REFERENCE, api.Q_UEL, api.Q_TANGENT_PRECEDENCE, api.Q_FORCE_LAST_IP])
def test_cohesive_element(orc, quirks):
    rng = np.random.default_rng(1)
    n = 64
    x0 = rng.uniform(-2, 2, (n, 1, 2))
    ang = rng.uniform(0, 2 * np.pi, n)
    L = rng.uniform(0.5, 2.0, n)
    tvec = np.stack([np.cos(ang), np.sin(ang)], 1) * L[:, None]
    x = np.concatenate([x0, x0 + tvec[:, None], x0 + tvec[:, None], x0], axis=1) + 1e-3 * rng.standard_normal((n, 4, 2))
    nodes = np.stack([rng.permutation(12)[:4] + 1 for _ in range(n)]).astype(np.int32)
    u = 4e-3 * rng.standard_normal(24)
    props = np.stack([rng.uniform(50, 200, n), rng.uniform(.005, .02, n), rng.uniform(.005, .02, n),
                      rng.uniform(.3, 1., n), rng.uniform(0., .5, n), rng.uniform(0, 1e-3, n)], 1)
    with api.Femx(quirks=quirks) as f:
        F, S = f.cohesive(x, nodes, u, props, 0.01)
    for e in range(n):
        Fo = orc.apply_tract_1(x[e], nodes[e], u, 0.01, props[e], quirks)
        So = orc.apply_stiff_1(x[e], nodes[e], u, 0.01, props[e], quirks)
        assert close(F[e], Fo, TOL_MAT), e
        assert close(S[e], So, TOL_MAT), e


def _meshes(golden):
    yield "lin_two_quads_qs", fm.read_inp(golden["decks"]["lin_two_quads_qs"])     # BASELINE config 1
    yield "two_quads_qs", fm.read_inp(golden["decks"]["two_quads_qs"])
    yield "cohesive_test_2", fm.read_inp(golden["decks"]["cohesive_test_2"])
    yield "hex6_jitter", mg.jitter(mg.hex_cube(6))
    yield "hex_rect", mg.hex_cube(5)
    yield "tet4_jitter", mg.jitter(mg.tet_cube(4))
    yield "quad_jitter", mg.jitter(mg.quad_rect(9, 7))
    yield "tri_jitter", mg.jitter(mg.tri_rect(9, 7))
    yield "coh_strip", mg.jitter(mg.cohesive_strip(12, 3))
    # a rank-local mesh: ragged node set, nl2g not the identity
    g = mg.jitter(mg.hex_cube(5))
    loc = fm.localize(g, mg.slab_partition(g, 3), 1)
    keep = np.isin(loc.bcnode[:, 0], loc.nl2g)       # BCs of nodes this rank holds
    yield "hex_rank1of3", dataclasses.replace(loc, bcnode=loc.bcnode[keep], bcval=loc.bcval[keep])


def test_pattern_values_fint_vs_oracle(fx, orc, golden):
    for name, m in _meshes(golden):
        fx.set_mesh(m).build_plan()
        om = orc.Mesh(m)
        rp_o, ci_o, va_o = orc.assemble_K(om, prealloc=m.pdim * (m.nodal_bw + 1)).export(om.ndof)
        rp, ci = fx.pattern()
        assert np.array_equal(rp, rp_o), name          # bit-exact pattern
        assert np.array_equal(ci, ci_o), name
        u = mg.displacement_field(m)
        vals, fint = fx.assemble_K_fint(u)
        assert close(vals, va_o, TOL_MAT), name
        assert np.array_equal(fx.assemble_K(), vals), name   # deterministic, same order
        assert close(fint, orc.fint(om, u.reshape(-1)), TOL_MAT), name
        # index map: slot(e,a,i,b,j) must address (row, col) of that entry
        off, bpos = fx.slot_map()
        for e in range(0, m.nels, max(1, m.nels // 40)):
            nd = m.elem_nodes(e)
            nen = len(nd)
            for a in range(nen):
                for b in range(nen):
                    row = m.pdim * (m.nl2g[nd[a] - 1] - 1)
                    col = m.pdim * (m.nl2g[nd[b] - 1] - 1)
                    slot = rp[row] + m.pdim * bpos[off[e] + a * nen + b]
                    assert ci[slot] == col, (name, e, a, b)
        # BC penalty: exact INSERT
        kb = fx.apply_kbc(vals.copy())
        _, _, vb_o = orc.assemble_K(om, apply_bc=True).export(om.ndof)
        pen = vb_o == float(np.float32(1e30))
        assert np.array_equal(kb[pen], vb_o[pen]) and close(kb[~pen], vb_o[~pen], TOL_MAT, np.abs(va_o).max()), name


def test_device_plan_equals_host_plan(golden):
    """The integer pass on the device (csrc/plan_gpu.cu: radix sorts, label propagation for the hex8 logical grid) against
    the host pass it replaces (csrc/plan.cpp, plan_hex.cpp; Femx(host_plan=True)): pattern, slot map and -- because every
    list the kernels read is identical -- the values of every assembly variant and the nodal stress, bit for bit."""
    cases = list(_meshes(golden)) + [("l_shaped_hex", _l_shaped(6)), ("hex_permuted", _permuted(mg.jitter(mg.hex_cube(6, clamp=False)), 3)),
                                     ("hex_box", mg.jitter(mg.hex_block(9, 5, 11))), ("tet_jitter", mg.jitter(mg.tet_cube(5))),
                                     ("tri_wide", mg.jitter(mg.tri_rect(40, 3))), ("hex_two_islands", _two_islands()),
                                     ("tet_shuffled", _shuffled(mg.jitter(mg.tet_cube(5)), 11)), ("quad_shuffled", _shuffled(mg.jitter(mg.quad_rect(12, 9)), 12))]
    for name, m in cases:
        u = mg.displacement_field(m)
        pure = set(np.unique(m.etype))
        variants = [0, 1, 9] + ([11] if pure <= {fm.TRI, fm.TET} else [])
        out = {}
        for host in (True, False):
            for av in variants:
                with api.Femx(assembly_variant=av, host_plan=host) as f:
                    f.set_mesh(m).build_plan()
                    rp, ci = f.pattern()
                    off, bpos = f.slot_map()
                    vals, fint = f.assemble_K_fint(u)
                    st = f.stress(u) if av == 0 else None
                    out[host, av] = (rp, ci, off, bpos, vals, fint, st)
        for av in variants:
            for x, y in zip(out[True, av], out[False, av]):
                assert (x is None and y is None) or np.array_equal(x, y), (name, av)
    # the hex8 tile kernels run on both plans (same logical grid, same row metadata)
    for name, m in (("hex6_jitter", mg.jitter(mg.hex_cube(6))), ("hex_permuted", _permuted(mg.jitter(mg.hex_cube(6, clamp=False)), 3)), ("l_shaped_hex", _l_shaped(6))):
        u = mg.displacement_field(m)
        for av in (12, 20):
            res = []
            for host in (True, False):
                with api.Femx(assembly_variant=av, host_plan=host) as f:
                    f.set_mesh(m).build_plan()
                    res.append(f.assemble_K_fint(u))
            assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1]), (name, av)


def _shuffled(m, seed):
    """Same mesh with random node numbering AND random element order (an unstructured mesh as far as the integer pass can tell)."""
    rng = np.random.default_rng(seed)
    perm = rng.permutation(m.nlnds)            # new id of old node i = perm[i]
    eo = rng.permutation(m.nels)
    nen = int(m.eptr[1] - m.eptr[0])
    conn = (perm[m.eind - 1] + 1).reshape(-1, nen)[eo]
    return dataclasses.replace(m, eind=conn.reshape(-1).astype(np.int32), coords=m.coords[np.argsort(perm)], emat=m.emat[eo], enlmat=m.enlmat[eo],
                               etype=m.etype[eo], bcnode=np.zeros((0, 1 + m.pdim), np.int32), bcval=np.zeros((0, m.pdim)))


def _two_islands():
    """Two disconnected hex8 blocks in one mesh: the logical-grid pass has to place the second component beside the first."""
    a, b = mg.jitter(mg.hex_block(3, 2, 2)), mg.jitter(mg.hex_block(2, 3, 2))
    nn = a.nnds
    eind = np.concatenate([a.eind, b.eind + nn]).astype(np.int32)
    eptr = np.concatenate([a.eptr, b.eptr[1:] + a.eptr[-1]]).astype(np.int32)
    coords = np.concatenate([a.coords, b.coords + np.array([10.0, 0.0, 0.0])])
    return dataclasses.replace(a, nnds=nn + b.nnds, eptr=eptr, eind=eind, etype=np.concatenate([a.etype, b.etype]),
                               emat=np.concatenate([a.emat, b.emat]), enlmat=np.concatenate([a.enlmat, b.enlmat]), coords=coords,
                               nl2g=np.arange(1, nn + b.nnds + 1, dtype=np.int32), bcnode=np.zeros((0, 4), np.int32), bcval=np.zeros((0, 3)))


def test_fused_rows_variant_vs_two_pass_and_oracle(orc, golden):
    """Variant 9 (fused rows kernel: element row blocks stay in shared memory) against variant 1 (two-pass through
    the packed Ke scratch) and the oracle, on every element mix incl. cohesive zeros and an unstructured hex mesh."""
    cases = list(_meshes(golden)) + [("l_shaped_hex", _l_shaped(6)), ("tet_jitter", mg.jitter(mg.tet_cube(5))),
                                     ("tri_wide", mg.jitter(mg.tri_rect(40, 3))), ("tet_rank1of2", None)]
    # three materials with different moduli AND Poisson ratios (rho = lambda / mu varies per element)
    three = np.array([[3e10, .25, 0, 0, 0], [7e9, .33, 0, 0, 0], [2.1e11, .1, 0, 0, 0]])
    for nm, base in (("tet_3mat", mg.jitter(mg.tet_cube(4))), ("tri_3mat", mg.jitter(mg.tri_rect(7, 6))), ("quad_3mat", mg.jitter(mg.quad_rect(8, 5)))):
        cases.append((nm, dataclasses.replace(base, mat=three, emat=(np.arange(base.nels) % 3 + 1).astype(np.int32))))
    for name, m in cases:
        if m is None:                                   # rank-local tet mesh: ragged node set, nl2g not the identity
            g = mg.jitter(mg.tet_cube(5))
            m = fm.localize(g, mg.slab_partition(g, 2), 1)
            m = dataclasses.replace(m, bcnode=m.bcnode[:0], bcval=m.bcval[:0])
        u = mg.displacement_field(m)
        om = orc.Mesh(m)
        _, _, va_o = orc.assemble_K(om).export(om.ndof)
        f_o = orc.fint(om, u.reshape(-1))
        res = {}
        simplex = bool(np.all((m.etype == fm.TRI) | (m.etype == fm.TET)))
        if not simplex:                                 # variant 11 (simplex rows kernel) must refuse other meshes
            with api.Femx(assembly_variant=11) as f:
                f.set_mesh(m).build_plan()
                with pytest.raises(api.FemxError):
                    f.assemble_K()
        for var in (9, 1) + ((11,) if simplex else ()):
            with api.Femx(assembly_variant=var) as f:
                f.set_mesh(m).build_plan()
                vals, fint = f.assemble_K_fint(u)
                assert close(vals, va_o, TOL_MAT), (name, var)
                assert close(fint, f_o, TOL_MAT), (name, var)
                assert np.array_equal(f.assemble_K(), vals), (name, var)
                res[var] = vals
        assert close(res[9], res[1], TOL_MAT), name      # same summation order except the split diagonal block
        if simplex:
            assert close(res[11], res[1], TOL_MAT), name


def test_inverted_simplex_falls_back_to_the_general_kernel(orc):
    """The simplex rows kernel scales gradients by sqrt(w detJ mu); a mesh with an inverted element (negative detJ, which
    the reference assembles as is) must be refused by variant 11 and routed to the fused rows kernel by auto."""
    m = mg.jitter(mg.tet_cube(3))
    eind = m.eind.copy()
    p = m.eptr[7]
    eind[p], eind[p + 1] = eind[p + 1], eind[p]
    m = dataclasses.replace(m, eind=eind)
    om = orc.Mesh(m)
    _, _, va_o = orc.assemble_K(om).export(om.ndof)
    u = mg.displacement_field(m)
    with api.Femx() as f:
        f.set_mesh(m).build_plan()
        vals, fint = f.assemble_K_fint(u)
        assert close(vals, va_o, TOL_MAT) and close(fint, orc.fint(om, u.reshape(-1)), TOL_MAT)
    with api.Femx(assembly_variant=11) as f:
        f.set_mesh(m).build_plan()
        with pytest.raises(api.FemxError):
            f.assemble_K()


def test_golden_matview_through_cuda(golden):
    """examples/tmp.txt: the Jacobian of cohesive_test_2.inp, now from the CUDA path."""
    g = golden["jacobian_matview_cohesive_test_2"]
    m = fm.read_inp(golden["decks"]["cohesive_test_2"])
    m = dataclasses.replace(m, mat=np.vstack([m.mat, np.zeros((1, 5))]), emat=np.array([1, 2, 0], np.int32))
    with api.Femx(quirks=api.Q_REFERENCE) as f:
        f.set_mesh(m).build_plan()
        rp, ci = f.pattern()
        K = f.apply_kbc(f.assemble_K())
        J = f.CalcJacobian(K, np.zeros(16), m.dt)
    for i in range(16):
        assert ci[rp[i]:rp[i + 1]].tolist() == [c for c, _ in g[i]]
    for i in range(8):
        for (c, gv), v in zip(g[i], J[rp[i]:rp[i + 1]]):
            assert abs(v - gv) <= 6e-6 * abs(gv) or (abs(gv) < 1e-300 and abs(v) < 1e-9), (i, c, v, gv)


@pytest.mark.parametrize("quirks", [api.Q_REFERENCE, api.Q_UEL])
def test_jacobian_residual_vs_oracle(orc, golden, quirks):
    for name, m in (("cohesive_test_2", fm.read_inp(golden["decks"]["cohesive_test_2"])),
                    ("two_quads_qs", fm.read_inp(golden["decks"]["two_quads_qs"])),
                    ("coh_strip", mg.jitter(mg.cohesive_strip(16, 3)))):
        om = orc.Mesh(m)
        rng = np.random.default_rng(4)
        u_tot = 2e-3 * rng.standard_normal(m.nlnds * 2)
        du = 1e-3 * rng.standard_normal(om.ndof)
        fext = 10.0 * rng.standard_normal(om.ndof)
        Ko = orc.assemble_K(om, apply_bc=True)
        Jo = orc.jacobian(om, Ko, u_tot, m.dt, quirks)
        _, _, jv_o = Jo.export(om.ndof)
        r_o = orc.residual(om, Jo, u_tot, du, fext, m.dt, m.t, quirks)
        with api.Femx(quirks=quirks) as f:
            f.set_mesh(m).build_plan()
            K = f.apply_kbc(f.assemble_K())
            J = f.CalcJacobian(K, u_tot, m.dt)
            r = f.CalcResidual(J, u_tot, du, fext, m.dt, m.t)
            br = f.residual_forces(u_tot, fext, m.dt, m.t)
        pen = jv_o == float(np.float32(1e30))
        scale = np.abs(jv_o[~pen]).max()
        assert np.array_equal(J[pen], jv_o[pen]), name
        assert close(J[~pen], jv_o[~pen], TOL_MAT, scale), name
        # rows with a 1e30 penalty: compare relative to the row's own magnitude
        assert close(r, r_o, TOL_MAT, None if np.abs(r_o).max() < 1e20 else np.abs(r_o[np.abs(r_o) < 1e20]).max()), name
        assert close(br, -(r_o - np.asarray(_spmv(Jo, du, om.ndof, orc))), 1e-9), name


def _spmv(Jo, x, n, orc):
    rp, ci, va = Jo.export(n)
    y = np.zeros(n)
    for i in range(n):
        y[i] = va[rp[i]:rp[i + 1]] @ x[ci[rp[i]:rp[i + 1]]]
    return y


def test_jacobian_update_equals_full_rebuild(golden):
    """femx_jacobian_update (Newton iterations after the first: only cohesive blocks and BC diagonals are rewritten)
    must give the full femx_jacobian bit for bit, for a sequence of different openings on a resident Jacobian."""
    import ctypes as C
    m = mg.jitter(mg.cohesive_strip(12, 3))
    nn = m.nlnds
    with api.Femx(quirks=api.Q_UEL) as f:
        f.set_mesh(m).build_plan()
        kv = f.assemble_K()
        d_k = f.device_buffer(8 * f.nnz).upload(kv)
        d_j = f.device_buffer(8 * f.nnz)
        rng = np.random.default_rng(5)
        for it in range(4):
            u = 1e-3 * rng.standard_normal((nn, 2))
            d_u = f.device_buffer(u.nbytes).upload(u)
            call = f.L.femx_jacobian if it == 0 else f.L.femx_jacobian_update
            f._ck(call(f.h, d_k.ptr, d_u.ptr, float(m.dt), d_j.ptr, api.DEVICE))
            assert np.array_equal(d_j.download(count=f.nnz), f.CalcJacobian(kv, u, m.dt)), it
        assert f.L.femx_jacobian_update(f.h, d_k.ptr, d_u.ptr, float(m.dt), d_k.ptr, api.DEVICE) != 0   # in place: refused


def test_form_rhs_vs_oracle(fx, orc, golden):
    """FormRHS on the device (nodal-force ramps, ApplyTraction + EdgeAreaNodes*, EnforceBCForce): the kernel adds each
    dof's contributions in the reference's call order with unfused roundings -> bit-exact against the oracle."""
    for name, m in _meshes(golden):
        fx.set_mesh(m).build_plan()
        om = orc.Mesh(m)
        fnode, fval, telsd, tval = mg.random_loads(m, seed=len(name))
        fx.set_loads(fnode, fval, telsd, tval)
        biggest = 0.0
        for t_init in (0.0, 0.5 * m.dt, 3.0 * m.dt):
            f = fx.FormRHS(t_init, m.dt, m.t)
            f_o = orc.form_rhs(om, fnode, fval, telsd, tval, t_init, m.dt, m.t)
            assert np.array_equal(f, f_o), (name, t_init, np.abs(f - f_o).max())
            biggest = max(biggest, np.abs(f_o).max())
        assert biggest > 0.0, name
    fx.set_loads()                                   # no loads: zeros + the BC override only
    f = fx.FormRHS(0.0, m.dt, m.t)
    assert np.array_equal(f, orc.form_rhs(om, [], [], [], [], 0.0, m.dt, m.t)), name


def test_stress_vs_oracle(fx, orc, golden):
    for name, m in _meshes(golden):
        fx.set_mesh(m).build_plan()
        u = mg.displacement_field(m)
        s_o, cnt = orc.stress(orc.Mesh(m), u.reshape(-1), orc.Q_REFERENCE)
        s = fx.stress(u)
        assert close(s, s_o, TOL_FIELD), name
        assert np.array_equal(fx.stress(u, layout=api.COL_MAJOR), s), name


def test_elastic_golden_vtk_through_cuda(fx, golden):
    coords = np.array([[0, 2], [0, 1], [0, 0], [1, 2], [1, 1], [1, 0]], float)
    m = fm.from_uniform(2, fm.QUA, [[1, 2, 5, 4], [2, 3, 6, 5]], coords, E=3640.0, nu=0.3)
    u = np.zeros((6, 2))
    u[:, 1] = 0.05 * coords[:, 1]
    s = fx.set_mesh(m).build_plan().stress(u)
    assert np.allclose(s, [105.0, 245.0, 0.0], rtol=1e-12, atol=1e-10)


def test_golden_vtk_files_regenerated_through_cuda(fx, golden):
    """The reference's two output files (examples/SNES_test/elastic) rebuilt from the CUDA stress recovery of the
    displacement they hold: identical text (WriteOutput, src/m_io.F90:480-555)."""
    for r in (0, 1):
        text = golden[f"elastic_vtk_rank{r}_text"]
        pts, cells, _, _, disp = fm.read_vtk(text)
        m = fm.from_uniform(2, fm.QUA, [[c + 1 for c in cell] for cell in cells], pts[:, :2], E=3640.0, nu=0.3)
        s = fx.set_mesh(m).build_plan().stress(disp[:, :2])
        unsign = lambda t: t.replace("-.000000", ".000000")   # sign of a shear that rounds to zero: solver round-off the file does not hold
        assert unsign(fm.write_vtk(m, s, disp[:, :2])) == unsign(text)


def test_edge_cases(fx):
    # ragged / mixed element types with unused global nodes, single element, empty BC list
    m = fm.from_uniform(2, fm.TRI, [[2, 5, 9]], np.array([[0, 0], [1, 0], [0, 1.]]))
    m = dataclasses.replace(m, nl2g=np.array([2, 5, 9], np.int32), nnds=11, eind=np.array([1, 2, 3], np.int32))
    fx.set_mesh(m).build_plan()
    rp, ci = fx.pattern()
    assert fx.nrows == 22 and fx.nnz == 36
    assert rp[2] == 0 and rp[4] == 12 and rp[-1] == 36
    assert ci[:6].tolist() == [2, 3, 8, 9, 16, 17]
    vals, fint = fx.assemble_K_fint(np.zeros(6))
    assert np.all(fint == 0) and abs(vals.reshape(6, 6).sum()) < 1e-3


def _fan(ntri):
    """ntri triangles around one hub node (more than 32 elements around a node: the row lists do not apply, two-pass path)"""
    ang = np.linspace(0.0, 2 * np.pi, ntri, endpoint=False)
    xy = np.concatenate([[[0.0, 0.0]], np.stack([np.cos(ang), np.sin(ang)], 1)])
    tris = [[1, 2 + i, 2 + (i + 1) % ntri] for i in range(ntri)]
    return fm.from_uniform(2, fm.TRI, tris, xy)


@pytest.mark.parametrize("host_plan", [False, True])
def test_plan_edge_cases(orc, host_plan):
    """Empty mesh, a local node no element touches, and a node with more than 32 elements around it, through the device
    integer pass and the host pass: same pattern, and values against the oracle where there are any."""
    # no elements, no nodes
    e = fm.from_uniform(2, fm.TRI, np.zeros((0, 3), np.int32), np.zeros((0, 2)))
    with api.Femx(host_plan=host_plan) as f:
        f.set_mesh(e).build_plan()
        rp, ci = f.pattern()
        assert f.nnz == 0 and len(ci) == 0 and np.all(rp == 0)
        assert len(f.assemble_K()) == 0
    # node 2 of 4 belongs to no element: its rows are empty
    m = fm.from_uniform(2, fm.TRI, [[1, 3, 4]], np.array([[0, 0], [5, 5], [1, 0], [0, 1.]]))
    with api.Femx(host_plan=host_plan) as f:
        f.set_mesh(m).build_plan()
        rp, ci = f.pattern()
        assert f.nnz == 36 and rp[2] == rp[3] == rp[4] == 12
        vals, fint = f.assemble_K_fint(np.zeros(8))
        om = orc.Mesh(m)
        rp_o, ci_o, va_o = orc.assemble_K(om).export(om.ndof)
        assert np.array_equal(rp, rp_o) and np.array_equal(ci, ci_o) and close(vals, va_o, TOL_MAT)
        s = f.stress(np.zeros(8))
        assert np.all(s[[0, 2, 3]] == 0) and not np.any(np.isfinite(s[1]))   # 0 / 0: the reference divides by count_node = 0 as well
    # 40 triangles around the hub
    fan = _fan(40)
    u = mg.displacement_field(fan)
    with api.Femx(host_plan=host_plan) as f:
        f.set_mesh(fan).build_plan()
        rp, ci = f.pattern()
        vals, fint = f.assemble_K_fint(u)
        om = orc.Mesh(fan)
        rp_o, ci_o, va_o = orc.assemble_K(om, prealloc=100).export(om.ndof)
        assert np.array_equal(rp, rp_o) and np.array_equal(ci, ci_o)
        assert close(vals, va_o, TOL_MAT) and close(fint, orc.fint(om, u.reshape(-1)), TOL_MAT)
        for av in (9, 11):
            with api.Femx(assembly_variant=av, host_plan=host_plan) as g:
                g.set_mesh(fan).build_plan()
                with pytest.raises(api.FemxError, match="32 elements"):
                    g.assemble_K()


def test_large_mesh_properties(fx):
    """Size-independent properties at a size the oracle cannot finish quickly: rigid-body
    modes give zero internal force, K is symmetric, row sums vanish, stress of a linear
    field is constant."""
    m = mg.jitter(mg.hex_cube(24))
    fx.set_mesh(m).build_plan()
    rp, ci = fx.pattern()
    x = m.coords
    rot = np.c_[-x[:, 1], x[:, 0], np.zeros(len(x))]
    vals, f_rot = fx.assemble_K_fint(rot)
    _, f_tr = fx.assemble_K_fint(np.tile([1.0, -2.0, 0.5], (len(x), 1)))
    scale = np.abs(vals).max()
    assert np.abs(f_tr).max() <= 1e-10 * scale and np.abs(f_rot).max() <= 1e-9 * scale * np.abs(x).max()
    import scipy.sparse as sp
    A = sp.csr_matrix((vals, ci, rp), shape=(fx.nrows, fx.nrows))
    assert abs(A - A.T).max() <= 1e-12 * scale
    G = 1e-3 * np.array([[1, 2, 0], [0, -1, 3], [2, 0, 1.0]])
    s = fx.stress(x @ G.T)
    assert np.abs(s - s[0]).max() <= 1e-9 * np.abs(s[0]).max()


def _permuted(m, seed):
    """Same mesh, random node numbering (exercises non-contiguous kmap runs)."""
    rng = np.random.default_rng(seed)
    perm = rng.permutation(m.nlnds)            # new id of old node i = perm[i]
    inv = np.argsort(perm)
    return dataclasses.replace(m, eind=(perm[m.eind - 1] + 1).astype(np.int32), coords=m.coords[inv],
                               bcnode=np.zeros((0, 4), np.int32), bcval=np.zeros((0, 3)))


def _l_shaped(n):
    """Hex cube with one octant of elements removed; unused nodes dropped and renumbered."""
    g = mg.jitter(mg.hex_cube(n, clamp=False))
    conn = g.eind.reshape(-1, 8)
    cen = g.coords[conn - 1].mean(1)
    keep = ~((cen[:, 0] > n / 2) & (cen[:, 1] > n / 2) & (cen[:, 2] > n / 2))
    conn = conn[keep]
    used = np.unique(conn)
    remap = np.zeros(g.nlnds + 1, np.int64)
    remap[used] = np.arange(1, len(used) + 1)
    return fm.from_uniform(3, fm.HEX, remap[conn], g.coords[used - 1])


def test_fused_hex_variant_vs_oracle_and_twopass(orc):
    g = mg.jitter(mg.hex_cube(7))
    cases = {
        "box_9x5x11": mg.hex_block(9, 5, 11),
        "cube17_chunks": mg.jitter(mg.hex_cube(17, clamp=False)),
        "box_20x30x4": mg.jitter(mg.hex_block(20, 30, 4)),
        "permuted": _permuted(mg.jitter(mg.hex_cube(6, clamp=False)), 3),
        "l_shaped": _l_shaped(8),
        "rank2of3": fm.localize(g, mg.slab_partition(g, 3), 2),
        "two_materials": dataclasses.replace(mg.jitter(mg.hex_cube(5, clamp=False)),
                                             mat=np.array([[3e10, .25, 0, 0, 0], [7e9, .25, 0, 0, 0]]),
                                             emat=(np.arange(125) % 2 + 1).astype(np.int32)),
    }
    for name, m in cases.items():
        m = dataclasses.replace(m, bcnode=np.zeros((0, 4), np.int32), bcval=np.zeros((0, 3)))
        u = mg.displacement_field(m)
        om = orc.Mesh(m)
        rp_o, ci_o, va_o = orc.assemble_K(om).export(om.ndof)
        f_o = orc.fint(om, u.reshape(-1))
        for var, lz in ((12, 0), (12, 25), (12, 3), (20, 0), (20, 25), (20, 3)):   # symmetric pair-block kernel, DMMA kernel
            with api.Femx(quirks=api.Q_REFERENCE, assembly_variant=var, lz=lz) as f4:
                f4.set_mesh(m).build_plan()
                rp, ci = f4.pattern()
                assert np.array_equal(rp, rp_o) and np.array_equal(ci, ci_o), name
                v4, f4int = f4.assemble_K_fint(u)
                assert close(v4, va_o, TOL_MAT), (name, var, lz)
                assert close(f4int, f_o, TOL_MAT), (name, var, lz)
                assert np.array_equal(f4.assemble_K(), v4), (name, var, lz)          # K-only kernel, same bits
                assert np.array_equal(f4.assemble_K_fint(u)[0], v4), (name, var, lz)  # deterministic: no atomics, fixed order
        with api.Femx(assembly_variant=1) as f1:
            v1 = f1.set_mesh(m).build_plan().assemble_K()
        assert close(v1, va_o, TOL_MAT), name


def test_material_change_after_the_plan(orc):
    """femx_set_materials after femx_build_plan: every factor the hex8 tile kernels derive from (E, nu) is recomputed, the
    kernels are re-validated (two Poisson ratios send auto to the general kernel), a smaller table is refused."""
    m = dataclasses.replace(mg.jitter(mg.hex_cube(6, clamp=False)), mat=np.array([[3e10, .25, 0, 0, 0], [7e9, .25, 0, 0, 0]]),
                            emat=(np.arange(216) % 2 + 1).astype(np.int32), bcnode=np.zeros((0, 4), np.int32), bcval=np.zeros((0, 3)))
    u = mg.displacement_field(m)
    for var in (0, 12, 20):
        with api.Femx(assembly_variant=var) as f:
            f.set_mesh(m).build_plan()
            v0, _ = f.assemble_K_fint(u)
            for mat in ([[1e9, .31, 0, 0, 0], [5e10, .31, 0, 0, 0]], [[2e10, .2, 0, 0, 0], [2e10, .35, 0, 0, 0]]):
                m2 = dataclasses.replace(m, mat=np.array(mat))
                om = orc.Mesh(m2)
                _, _, va_o = orc.assemble_K(om).export(om.ndof)
                f.set_materials(m2.mat)
                if var != 0 and mat[0][1] != mat[1][1]:
                    with pytest.raises(api.FemxError, match="Poisson"):
                        f.assemble_K_fint(u)
                    continue
                v, fi = f.assemble_K_fint(u)
                assert close(v, va_o, TOL_MAT), (var, mat)
                assert close(fi, orc.fint(om, u.reshape(-1)), TOL_MAT), (var, mat)
            with pytest.raises(api.FemxError, match="material id"):
                f.set_materials(m.mat[:1])
    mt = dataclasses.replace(mg.jitter(mg.tet_cube(3)), bcnode=np.zeros((0, 4), np.int32), bcval=np.zeros((0, 3)))
    with api.Femx() as f:                                        # simplex rows kernel: same refresh
        f.set_mesh(mt).build_plan()
        f.assemble_K()
        m2 = dataclasses.replace(mt, mat=np.array([[1.7e9, .4, 0, 0, 0]]))
        om = orc.Mesh(m2)
        f.set_materials(m2.mat)
        assert close(f.assemble_K(), orc.assemble_K(om).export(om.ndof)[2], TOL_MAT)


def test_mirrored_or_inverted_hex_falls_back_to_the_general_kernel(orc):
    """The hex8 tile kernels scale gradients by rsqrt(detJ).  A left-handed grid (every detJ < 0) passes the orientation
    walk; the reference assembles such elements as they are (src/m_local.F90:61-65), so auto must use the general kernel
    and an explicit request for a tile kernel must be refused."""
    g = mg.jitter(mg.hex_cube(4, clamp=False))
    mirrored = dataclasses.replace(g, coords=g.coords * np.array([1.0, 1.0, -1.0]))
    conn = g.eind.reshape(-1, 8).copy()
    bad = g.coords.copy()
    bad[conn[21, 6] - 1] = bad[conn[21, 0] - 1] - 0.4          # one node pushed through the opposite corner: detJ < 0 at some points
    for m in (mirrored, dataclasses.replace(g, coords=bad)):
        m = dataclasses.replace(m, bcnode=np.zeros((0, 4), np.int32), bcval=np.zeros((0, 3)))
        om = orc.Mesh(m)
        _, _, va_o = orc.assemble_K(om).export(om.ndof)
        u = mg.displacement_field(m)
        with api.Femx() as f:
            f.set_mesh(m).build_plan()
            vals, fint = f.assemble_K_fint(u)
            assert np.isfinite(vals).all()
            assert close(vals, va_o, TOL_MAT) and close(fint, orc.fint(om, u.reshape(-1)), TOL_MAT)
        for var in (12, 20):
            with api.Femx(assembly_variant=var) as f:
                f.set_mesh(m)
                with pytest.raises(api.FemxError, match="detJ"):
                    f.build_plan()


def test_hex_tile_variants_refuse_unqualified_mesh():
    with pytest.raises(api.FemxError, match="assembly_variant"):
        api.Femx(assembly_variant=2)               # the measured dead ends of round 1 are gone from the library
    m = mg.tet_cube(2)
    with api.Femx(assembly_variant=20) as f:
        f.set_mesh(m)
        with pytest.raises(api.FemxError, match="does not qualify"):
            f.build_plan()
    m2 = dataclasses.replace(mg.hex_cube(3, clamp=False), mat=np.array([[3e10, .25, 0, 0, 0], [3e10, .3, 0, 0, 0]]),
                             emat=(np.arange(27) % 2 + 1).astype(np.int32))
    with api.Femx(assembly_variant=12) as f:
        f.set_mesh(m2)
        with pytest.raises(api.FemxError, match="Poisson"):
            f.build_plan()
    with api.Femx() as f:                          # a second plan on the same ctx releases the first one's device arrays
        for n in (5, 3, 6):
            mm = mg.hex_cube(n)
            assert f.set_mesh(mm).build_plan().assemble_K().shape[0] == f.nnz
    with api.Femx(assembly_variant=0) as f:        # auto: falls back to the two-pass path
        assert f.set_mesh(m2).build_plan().assemble_K().shape[0] == f.nnz
