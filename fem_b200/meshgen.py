"""Synthetic meshes of BASELINE.json's configs, in the reference's conventions.

Node numbering is lexicographic, x fastest, unit spacing (make_example.F90:64-85); default
material E=3.0e10, nu=0.25 (make_example.F90:91).  Every generator returns a `Mesh` whose
arrays go straight into the C ABI, and `mesh.write_inp` emits the current `.inp` grammar.

  hex_cube(n)        C2 / C5: n^3 hex8, node order of ShapeFuncdHex (src/m_elems.F90:574-579)
  tet_cube(m)        C3: Kuhn 6-tet split of an m^3 cube, positive detJ
  quad_rect(nx, ny)  plane-strain quad4 sheet
  cohesive_strip(nx, ny)  C4: two quad4 blocks joined by nx cohesive elements
                     (cohesive_test_2.inp scaled; node order of Cohesive_element.f:58-60)
"""
from __future__ import annotations

import numpy as np

from .mesh import TRI, COH, HEX, QUA, TET, Mesh, from_uniform

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = x.astype(np.uint64) + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def hash_uniform(ids: np.ndarray, seed: int, ncomp: int) -> np.ndarray:
    """U(-1, 1) per (id, component) from splitmix64(seed ^ id) chained per component."""
    out = np.empty((len(ids), ncomp))
    s = splitmix64(np.uint64(seed) ^ ids.astype(np.uint64))
    for c in range(ncomp):
        s = splitmix64(s)
        out[:, c] = (s >> np.uint64(11)).astype(np.float64) * (2.0 / (1 << 53)) - 1.0
    return out


def jitter(m: Mesh, amount=0.15, seed=12345, lo=None, hi=None) -> Mesh:
    """Move interior nodes by amount*h*U(-1,1) so elements are not congruent.  The offset of a
    node depends only on its GLOBAL id, so rank-local pieces of one mesh agree (pass the global
    bounding box as lo/hi for those)."""
    lo = m.coords.min(0) if lo is None else np.asarray(lo, float)
    hi = m.coords.max(0) if hi is None else np.asarray(hi, float)
    interior = np.all((m.coords > lo + 1e-9) & (m.coords < hi - 1e-9), axis=1)
    d = amount * hash_uniform(m.nl2g.astype(np.uint64), seed, m.pdim)
    c = m.coords.copy()
    c[interior] += d[interior]
    import dataclasses
    return dataclasses.replace(m, coords=c)


def displacement_field(m: Mesh, seed=20261104) -> np.ndarray:
    """u(x) = G x + 1e-3 sin(.) + 1e-4 noise, [nlnds, pdim] in local node order."""
    pd = m.pdim
    G = 1e-3 * np.array([[0.7, -0.3, 0.2], [0.4, 0.9, -0.5], [-0.2, 0.1, 0.6]])[:pd, :pd]
    x = m.coords
    u = x @ G.T
    u += 1e-3 * np.sin(0.37 * x[:, ::-1] + 0.11 * np.arange(pd))
    u += 1e-4 * hash_uniform(m.nl2g.astype(np.uint64), seed, pd)
    return np.ascontiguousarray(u)


def _grid_ids(shape):
    """1-based node ids on a lexicographic grid, x fastest; returns an array indexed [i, j(, k)]."""
    n = int(np.prod(shape))
    return (np.arange(1, n + 1, dtype=np.int64)).reshape(shape[::-1]).transpose(*range(len(shape) - 1, -1, -1))


def hex_cube(n: int, E=3.0e10, nu=0.25, clamp=True) -> Mesh:
    nx = ny = nz = n
    ids = _grid_ids((nx + 1, ny + 1, nz + 1))
    i, j, k = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    # element order: x fastest as well
    order = np.argsort((i + nx * (j + ny * k)).reshape(-1), kind="stable")
    i, j, k = (a.reshape(-1)[order] for a in (i, j, k))
    conn = np.stack([ids[i, j, k], ids[i + 1, j, k], ids[i + 1, j + 1, k], ids[i, j + 1, k],
                     ids[i, j, k + 1], ids[i + 1, j, k + 1], ids[i + 1, j + 1, k + 1], ids[i, j + 1, k + 1]], axis=1)
    zz, yy, xx = np.meshgrid(np.arange(nz + 1.0), np.arange(ny + 1.0), np.arange(nx + 1.0), indexing="ij")
    coords = np.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], axis=1)
    m = from_uniform(3, HEX, conn, coords, E=E, nu=nu, nodal_bw=26)
    if clamp:  # clamp face x=0
        face = ids[0, :, :].reshape(-1)
        bc = np.zeros((len(face), 4), np.int32)
        bc[:, 0] = np.sort(face)
        m.bcnode, m.bcval = bc, np.zeros((len(face), 3))
    return m


_KUHN = [((0, 1, 2), 1), ((0, 2, 1), -1), ((1, 0, 2), -1), ((1, 2, 0), 1), ((2, 0, 1), 1), ((2, 1, 0), -1)]


def tet_cube(mdiv: int, E=3.0e10, nu=0.25) -> Mesh:
    n = mdiv
    ids = _grid_ids((n + 1, n + 1, n + 1))
    i, j, k = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
    order = np.argsort((i + n * (j + n * k)).reshape(-1), kind="stable")
    base = np.stack([a.reshape(-1)[order] for a in (i, j, k)], axis=1)  # [ncell, 3]
    tets = []
    for perm, sign in _KUHN:
        p = [base.copy()]
        cur = base.copy()
        for ax in perm:
            cur = cur.copy()
            cur[:, ax] += 1
            p.append(cur)
        v = [ids[q[:, 0], q[:, 1], q[:, 2]] for q in p]
        if sign < 0:
            v[1], v[2] = v[2], v[1]
        tets.append(np.stack(v, axis=1))
    conn = np.stack(tets, axis=1).reshape(-1, 4)  # 6 tets per cell, cells x-fastest
    zz, yy, xx = np.meshgrid(np.arange(n + 1.0), np.arange(n + 1.0), np.arange(n + 1.0), indexing="ij")
    coords = np.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], axis=1)
    return from_uniform(3, TET, conn, coords, E=E, nu=nu, nodal_bw=26)


def quad_rect(nx: int, ny: int, E=3.0e10, nu=0.25) -> Mesh:
    ids = _grid_ids((nx + 1, ny + 1))
    i, j = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    order = np.argsort((i + nx * j).reshape(-1), kind="stable")
    i, j = i.reshape(-1)[order], j.reshape(-1)[order]
    conn = np.stack([ids[i, j], ids[i + 1, j], ids[i + 1, j + 1], ids[i, j + 1]], axis=1)
    yy, xx = np.meshgrid(np.arange(ny + 1.0), np.arange(nx + 1.0), indexing="ij")
    coords = np.stack([xx.reshape(-1), yy.reshape(-1)], axis=1)
    return from_uniform(2, QUA, conn, coords, E=E, nu=nu, nodal_bw=8)


def tri_rect(nx: int, ny: int, E=3.0e10, nu=0.25) -> Mesh:
    """nx x ny unit cells cut into two tri3 each (alternating diagonals, so nodes see 4 to 8 triangles)."""
    q = quad_rect(nx, ny, E=E, nu=nu)
    c = q.eind.reshape(-1, 4)
    flip = ((np.arange(len(c)) % nx + np.arange(len(c)) // nx) % 2).astype(bool)
    t1 = np.where(flip[:, None], c[:, [0, 1, 3]], c[:, [0, 1, 2]])
    t2 = np.where(flip[:, None], c[:, [1, 2, 3]], c[:, [0, 2, 3]])
    conn = np.stack([t1, t2], axis=1).reshape(-1, 3)
    return from_uniform(2, TRI, conn, q.coords, E=E, nu=nu, nodal_bw=8)


def cohesive_strip(nx: int, ny: int, E=3640.0, nu=0.3, props=(100.0, 0.01, 0.01, 1.0, 0.0, 0.0),
                   t=0.02, dt=0.01) -> Mesh:
    """Two nx x ny quad4 blocks stacked in y with duplicated interface nodes and nx cohesive
    elements between them.  Bottom clamped; top u_y prescribed; all u_x fixed on those rows
    (the scaled-up BCs of examples/cohesive_test_2.inp:25-28)."""
    bot = quad_rect(nx, ny, E, nu)
    nnb = bot.nlnds
    conn_b = bot.eind.reshape(-1, 4)
    conn_t = conn_b + nnb
    coords = np.vstack([bot.coords, bot.coords + np.array([0.0, float(ny)])])
    ids_b = _grid_ids((nx + 1, ny + 1))
    i = np.arange(nx)
    coh = np.stack([ids_b[i, ny], ids_b[i + 1, ny], ids_b[i + 1, 0] + nnb, ids_b[i, 0] + nnb], axis=1)
    nq = 2 * nx * ny
    etype = np.concatenate([np.full(nq, QUA), np.full(nx, COH)]).astype(np.int32)
    eind = np.concatenate([conn_b.reshape(-1), conn_t.reshape(-1), coh.reshape(-1)]).astype(np.int32)
    eptr = np.arange(0, 4 * (nq + nx) + 1, 4, dtype=np.int32)
    emat = np.concatenate([np.ones(nq), np.zeros(nx)]).astype(np.int32)
    enl = np.concatenate([np.zeros(nq), np.ones(nx)]).astype(np.int32)
    bottom = ids_b[:, 0]
    top = ids_b[:, ny] + nnb
    bcnode = np.zeros((2 * (nx + 1), 3), np.int32)
    bcnode[:, 0] = np.concatenate([bottom, top])
    bcval = np.zeros((2 * (nx + 1), 2))
    bcval[nx + 1:, 1] = 0.1
    return Mesh(pdim=2, etype=etype, eptr=eptr, eind=eind, emat=emat, enlmat=enl, coords=coords,
                nl2g=np.arange(1, len(coords) + 1), nnds=len(coords), mat=np.array([[E, nu, 1e18, 1.0, 3000.0]]),
                cohprops=np.array([props]), bcnode=bcnode, bcval=bcval, t=t, dt=dt, nodal_bw=10).normalized()


def hex_block(nx: int, ny: int, nz: int, z0: int = 0, nz_total: int | None = None, E=3.0e10, nu=0.25,
              jittered=True) -> Mesh:
    """Rank-local z-slab [z0, z0+nz] of an nx x ny x nz_total hex8 column, built directly in
    local form (what `localize` would give for a slab partition): global node ids of the
    column, local ids ascending with them, elements in ascending global order."""
    nz_total = nz if nz_total is None else nz_total
    npl = (nx + 1) * (ny + 1)
    i, j, k = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    order = np.argsort((i + nx * (j + ny * k)).reshape(-1), kind="stable")
    i, j, k = (a.reshape(-1)[order] for a in (i, j, k))
    nid = lambda a, b, c: 1 + a + (nx + 1) * (b + (ny + 1) * c)       # local == global - z0*npl
    conn = np.stack([nid(i, j, k), nid(i + 1, j, k), nid(i + 1, j + 1, k), nid(i, j + 1, k),
                     nid(i, j, k + 1), nid(i + 1, j, k + 1), nid(i + 1, j + 1, k + 1), nid(i, j + 1, k + 1)], axis=1)
    zz, yy, xx = np.meshgrid(np.arange(z0, z0 + nz + 1.0), np.arange(ny + 1.0), np.arange(nx + 1.0), indexing="ij")
    coords = np.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], axis=1)
    m = from_uniform(3, HEX, conn, coords, E=E, nu=nu, nodal_bw=26)
    m.nl2g = (np.arange(1, len(coords) + 1) + z0 * npl).astype(np.int32)
    m.nnds = npl * (nz_total + 1)
    m.eglobal = np.arange(m.nels, dtype=np.int64) + z0 * nx * ny
    if jittered:
        m = jitter(m, lo=(0, 0, 0), hi=(nx, ny, nz_total))
    return m


def random_loads(m: Mesh, nfrcs=40, ntrcs=60, seed=7):
    """Synthetic FormRHS input (src/m_global.F90:373-436): nodal forces on GLOBAL nodes this rank holds (repeats
    included, so several contributions meet in one dof) and tractions on random sides of random non-cohesive LOCAL
    elements, with ramp windows that start before, inside and after [t_init, t_init + dt]; one traction row with
    element 0 (skipped by the reference, :427).  Returns fnode, fval, telsd, tval."""
    rng = np.random.default_rng(seed)
    pd = m.pdim
    fnode = rng.choice(m.nl2g, size=nfrcs).astype(np.int32)
    fval = np.c_[rng.uniform(-1e3, 1e3, (nfrcs, pd)), rng.uniform(0.0, 0.6, nfrcs), rng.uniform(0.7, 2.0, nfrcs)]
    solid = np.flatnonzero(np.asarray(m.etype) != 5)
    els = rng.choice(solid, size=ntrcs)
    nsides = np.where(m.etype[els] == 1, 3, np.where(m.etype[els] == 4, 6, 4))
    telsd = np.c_[els + 1, 1 + rng.integers(0, 1 << 30, ntrcs) % nsides].astype(np.int32)
    telsd[ntrcs // 2, 0] = 0
    tval = np.c_[rng.uniform(-50.0, 50.0, (ntrcs, pd)), rng.uniform(0.0, 0.6, ntrcs) * m.dt, rng.uniform(0.7, 2.0, ntrcs) * m.dt]
    return fnode, fval, telsd, tval


def slab_partition(m: Mesh, nparts: int) -> np.ndarray:
    """Element partition into `nparts` contiguous element ranges (slabs along the slowest
    axis for the structured generators).  Stand-in for METIS_PartMeshNodal
    (src/m_io.F90:137): the hot path only consumes `epart`."""
    return (np.arange(m.nels, dtype=np.int64) * nparts // max(m.nels, 1)).astype(np.int32)


def block_grid(nranks: int):
    """(Px, Py, Pz) of the block partition bench.py uses at N ranks: 2x1x1, 2x2x1, 2x2x2 (BASELINE configs[4] is the last)."""
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[nranks]


def hex_block_rank(n: int, grid, rank: int, E=3.0e10, nu=0.25, jittered=True) -> Mesh:
    """Rank-local mesh of block `rank` of a (Px n) x (Py n) x (Pz n) hex8 cube cut into Px x Py x Pz blocks of n^3 elements
    (BASELINE configs[4] at grid (2, 2, 2), n = 100), built without ever forming the global mesh.  Global node numbering is
    contiguous per block: a node on a shared face belongs to the lower block, blocks are numbered x-fastest,
    nodes inside a block lexicographically.  Local ids ascend with the global id (src/main.F90:57-63), elements are in
    lexicographic order, node order of an element as ShapeFuncdHex (src/m_elems.F90:574-579).  Interior nodes of the WHOLE
    cube are jittered by a hash of the global id, so the pieces of different ranks agree on shared nodes."""
    Px, Py, Pz = grid
    bx, by, bz = rank % Px, (rank // Px) % Py, rank // (Px * Py)
    n1 = n + 1
    own_len = lambda b: n1 if b == 0 else n                     # owned node positions of a block along one axis
    counts = np.array([[[own_len(x) * own_len(y) * own_len(z) for x in range(Px)] for y in range(Py)] for z in range(Pz)], np.int64)
    offs = np.concatenate([[0], np.cumsum(counts.reshape(-1))])     # block order: x fastest
    i = np.arange(n1)

    def axis(b):        # owner block coordinate and position inside the owner's owned range, per lattice position
        ob = np.where((i == 0) & (b > 0), b - 1, b)
        pos = np.where((i == 0) & (b > 0), n, i)                # position in the owner's lattice
        pos = np.where(ob == 0, pos, pos - 1)                   # owned range starts at lattice position 1 unless the block is first
        return ob, pos

    obx, px = axis(bx); oby, py = axis(by); obz, pz = axis(bz)
    OBX, OBY, OBZ = np.meshgrid(obx, oby, obz, indexing="ij")   # [i, j, k]
    PX, PY, PZ = np.meshgrid(px, py, pz, indexing="ij")
    lx = np.where(OBX == 0, n1, n); ly = np.where(OBY == 0, n1, n)
    gid = offs[(OBZ * Py + OBY) * Px + OBX] + (PZ * ly + PY) * lx + PX + 1      # 1-based global node id, [i, j, k]
    gflat = gid.transpose(2, 1, 0).reshape(-1)                  # lattice order: i fastest
    order = np.argsort(gflat, kind="stable")
    nl2g = gflat[order]
    local_of = np.empty(len(order), np.int64)
    local_of[order] = np.arange(1, len(order) + 1)
    X, Y, Z = np.meshgrid(bx * n + i, by * n + i, bz * n + i, indexing="ij")
    coords = np.stack([X.transpose(2, 1, 0).reshape(-1), Y.transpose(2, 1, 0).reshape(-1), Z.transpose(2, 1, 0).reshape(-1)], 1).astype(float)
    lid = lambda a, b, c: local_of[(c * n1 + b) * n1 + a]
    e = np.arange(n)
    EI, EJ, EK = np.meshgrid(e, e, e, indexing="ij")
    EI, EJ, EK = (a.transpose(2, 1, 0).reshape(-1) for a in (EI, EJ, EK))
    conn = np.stack([lid(EI, EJ, EK), lid(EI + 1, EJ, EK), lid(EI + 1, EJ + 1, EK), lid(EI, EJ + 1, EK),
                     lid(EI, EJ, EK + 1), lid(EI + 1, EJ, EK + 1), lid(EI + 1, EJ + 1, EK + 1), lid(EI, EJ + 1, EK + 1)], 1)
    from .mesh import from_uniform
    m = from_uniform(3, HEX, conn, coords[order], E=E, nu=nu)
    import dataclasses
    m = dataclasses.replace(m, nl2g=nl2g.astype(np.int32), nnds=int(offs[-1]), bcnode=np.zeros((0, 4), np.int32), bcval=np.zeros((0, 3))).normalized()
    if jittered:
        m = jitter(m, lo=(0, 0, 0), hi=(Px * n, Py * n, Pz * n))
    return m
