"""Multi-rank parity through the C-ABI.

* ONE GPU, 2 and 3 ranks (no skip): the ranks are femx contexts of this process on device 0 (femx_peer_connect_local).  The
  whole data path of csrc/peer.cu runs -- per-owner value ranges, arena layout, remote stores of the push kernel, step flags,
  the O(interface) receiver add -- and the owned rows / internal force are compared with the 1-rank oracle.  Kernels of
  such ranks run one after another on the device, so none may wait for a peer: assemblies signal only and
  femx_peer_complete finishes every rank afterwards (spinning on a rank that has not been launched is exactly what the
  B200 profiling notes forbid on a shared GPU).
* rank-local meshes with cohesive elements on the partition interface: femx_jacobian, femx_residual_forces and
  femx_form_rhs per rank, partial results summed over the ranks, against the 1-rank oracle (R14, src/m_global.F90:98-235).
* tools/mgpu_parity.py under torchrun (two processes; on a 1-GPU box both on GPU 0, deferred): one process per GPU, IPC-mapped
  arenas and the in-kernel step barrier among neighbours.
"""
import dataclasses
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from fem_b200 import api, mesh as fm, meshgen as mg, parallel as par

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-12


def _close(a, b, scale):
    return bool(np.all(np.abs(a - b) <= TOL * np.maximum(np.abs(b), scale)))


def _nobc(m):
    return dataclasses.replace(m, bcnode=np.zeros((0, 1 + m.pdim), np.int32), bcval=np.zeros((0, m.pdim)))


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("kind,variant", [("hex", 0), ("hex", 12), ("tet", 0), ("quad", 0)])
def test_local_ranks_exchange_on_one_gpu(orc, world, kind, variant):
    g = {"hex": lambda: mg.jitter(mg.hex_cube(9, clamp=False)), "tet": lambda: mg.jitter(mg.tet_cube(5)),
         "quad": lambda: mg.jitter(mg.quad_rect(23, 17))}[kind]()
    g = _nobc(g)
    epart = mg.slab_partition(g, world)
    og = orc.Mesh(g)
    rp_g, ci_g, va_g = orc.assemble_K(og).export(og.ndof)
    ug = mg.displacement_field(g)
    f_ref = orc.fint(og, ug.reshape(-1))
    fxs, locs = [], []
    try:
        for r in range(world):
            loc = fm.localize(g, epart, r)
            fx = api.Femx(assembly_variant=variant)
            fx.set_mesh(loc).build_plan()
            fxs.append(fx); locs.append(loc)
        lr = par.LocalRanks(fxs)
        assert sum(int(ep.recv_counts.sum()) for ep in lr.eps) > 0
        for rep in range(3):                                   # both arena buffers, then the first again
            owned, f_tot = lr.assemble_K_fint([mg.displacement_field(loc) for loc in locs])
            for r, (rp, ci, va) in enumerate(owned):
                r0, r1 = lr.eps[r].row_start, lr.eps[r].row_end
                assert (r0, r1) == fm.petsc_ownership(og.ndof, world, r)
                assert np.array_equal(rp, rp_g[r0:r1 + 1] - rp_g[r0]), (r, rep)
                assert np.array_equal(ci, ci_g[rp_g[r0]:rp_g[r1]]), (r, rep)
                assert _close(va, va_g[rp_g[r0]:rp_g[r1]], np.abs(va_g).max()), (r, rep)
                assert len(lr.eps[r].if_rows) < (r1 - r0)      # the receiver add touches interface rows only
            assert _close(f_tot, f_ref, np.abs(f_ref).max()), rep
            # GetVec_U (R17): owners push the ghost displacements; every rank's local vector must be the global one gathered
            ugl = ug.reshape(-1) + rep
            for r, ul in enumerate(lr.scatter_u(ugl)):
                assert np.array_equal(ul, ugl.reshape(-1, g.pdim)[locs[r].nl2g - 1]), (r, rep)
        lr.close()
    finally:
        for fx in fxs:
            fx.close()


@pytest.mark.parametrize("world", [2, 3])
def test_jacobian_residual_rhs_on_rank_local_meshes(orc, world):
    """Cohesive strip cut so that the cohesive elements, the block below and the block above sit on different ranks."""
    g = mg.cohesive_strip(9, 4)
    fnode, fval, telsd, tval = mg.random_loads(g, nfrcs=15, ntrcs=20)
    g = _nobc(dataclasses.replace(g, fnode=fnode, fval=fval, telsd=telsd, tval=tval))
    epart = mg.slab_partition(g, world)
    assert len(set(epart[np.asarray(g.etype) == fm.COH])) >= 1
    og = orc.Mesh(g)
    rng = np.random.default_rng(3)
    ug = 2e-3 * rng.standard_normal((g.nnds, 2))
    ug[g.nnds // 2:, 1] += 4e-3                                # open the interface
    t_init, dt, t_tot = 0.0, g.dt, g.t
    K_g = orc.assemble_K(og)
    J_g = orc.jacobian(og, K_g, ug.reshape(-1), dt, api.Q_UEL)
    rp_g, ci_g, vj_g = J_g.export(og.ndof)
    f_g = orc.form_rhs(og, g.fnode, g.fval, g.telsd, g.tval, t_init, dt, t_tot)
    br_g = -orc.residual(og, J_g, ug.reshape(-1), np.zeros(og.ndof), f_g, dt, t_tot, api.Q_UEL)      # R = J du - bracket, du = 0
    n = og.ndof
    J_sum, f_sum, br_sum = np.zeros((n, n)), np.zeros(n), np.zeros(n)
    for r in range(world):
        loc = fm.localize(g, epart, r)
        with api.Femx(quirks=api.Q_UEL) as f:
            f.set_mesh(loc).build_plan()
            rp, ci = f.pattern()
            ul = ug[loc.nl2g - 1]
            J = f.CalcJacobian(f.assemble_K(), ul, dt)
            rows = np.repeat(np.arange(n), np.diff(rp))
            np.add.at(J_sum, (rows, ci), J)
            f.set_loads(loc.fnode, loc.fval, loc.telsd, loc.tval)
            fr = f.FormRHS(t_init, dt, t_tot)
            f_sum += fr
            br_sum += f.residual_forces(ul, fr, dt, t_tot)
    J_ref = np.zeros((n, n))
    J_ref[np.repeat(np.arange(n), np.diff(rp_g)), ci_g] = vj_g
    assert _close(J_sum, J_ref, np.abs(vj_g).max())
    assert _close(f_sum, f_g, np.abs(f_g).max())
    assert _close(br_sum, br_g, np.abs(br_g).max())


@pytest.mark.parametrize("renumber", [False, True])
def test_local_ranks_on_a_metis_partition(orc, renumber):
    """METIS_PartMeshNodal (called as src/m_io.F90:137 calls it) into 3 parts, through the device exchange: with the mesh
    file's numbering and the PETSC_DECIDE split (what the reference does: most rows a rank assembles belong to somebody else),
    and with the nodes renumbered part by part and the parts' own row ranges (interface rows only)."""
    world = 3
    g = _nobc(mg.jitter(mg.tet_cube(6)))
    try:
        epart, cut = fm.metis_partition(g, world)
    except RuntimeError as e:
        pytest.skip(str(e))
    row_bounds = None
    if renumber:
        g, nb = fm.renumber_by_partition(g, epart, world)
        row_bounds = g.pdim * nb
    og = orc.Mesh(g)
    rp_g, ci_g, va_g = orc.assemble_K(og).export(og.ndof)
    f_ref = orc.fint(og, mg.displacement_field(g).reshape(-1))
    fxs, locs = [], []
    try:
        for r in range(world):
            loc = fm.localize(g, epart, r)
            fx = api.Femx()
            fx.set_mesh(loc).build_plan()
            fxs.append(fx); locs.append(loc)
        lr = par.LocalRanks(fxs, row_bounds)
        sent = sum(int(ep.recv_counts.sum()) for ep in lr.eps)
        for rep in range(2):
            owned, f_tot = lr.assemble_K_fint([mg.displacement_field(loc) for loc in locs])
            for r, (rp, ci, va) in enumerate(owned):
                r0, r1 = lr.eps[r].row_start, lr.eps[r].row_end
                assert np.array_equal(rp, rp_g[r0:r1 + 1] - rp_g[r0]) and np.array_equal(ci, ci_g[rp_g[r0]:rp_g[r1]]), (r, rep)
                assert _close(va, va_g[rp_g[r0]:rp_g[r1]], np.abs(va_g).max()), (r, rep)
            assert _close(f_tot, f_ref, np.abs(f_ref).max()), rep
        if renumber:   # only interface rows cross: a small fraction of the matrix
            assert all(len(ep.if_rows) < (ep.row_end - ep.row_start) for ep in lr.eps) and sent < 0.5 * len(va_g)
        lr.close()
    finally:
        for fx in fxs:
            fx.close()


def test_two_rank_parity_all_exchanges():
    """two PROCESSES under torchrun, arenas mapped through CUDA IPC: one per GPU where the box has two (NCCL + both peer modes,
    device-side step barrier), else both on GPU 0 (gloo setup, deferred completion -- the more-ranks-than-GPUs layout)"""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "mgpu_parity.py")]
    env = dict(os.environ)
    if torch.cuda.device_count() < 2:
        env["FEMX_SAME_DEVICE"] = "1"
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MGPU PARITY OK" in r.stdout and "MISMATCH" not in r.stdout
