"""Mesh container, `.inp` reader/writer and rank-local renumbering.

Host-side mirror of the reference's input layer -- same grammar, same index
conventions -- so a deck the reference reads gives the arrays the C ABI takes:

* grammar: `ReadParameters` src/m_io.F90:12-20, `ReadElementsCoords` :71-105,
  `ReadDistMaterials` :282-328, forces :330-371, BCs :373-412;
* local numbering: src/main.F90:49-96 (`work`/`nl2g`/`indxmap`);
* element distribution: `epart` -> ascending global element order per rank,
  src/m_io.F90:200-237.

Element type codes are `getElTypeNo` (src/m_elems.F90:25): 1 tri, 2 qua, 3 tet,
4 hex, 5 coh.  hex has 8 nodes here (the reference's table says 6: defect D1).
"""
from __future__ import annotations

import dataclasses
from typing import Optional

import numpy as np

TRI, QUA, TET, HEX, COH = 1, 2, 3, 4, 5
TYPE_NAMES = {"tri": TRI, "qua": QUA, "tet": TET, "hex": HEX, "coh": COH}
TYPE_TAGS = {v: k for k, v in TYPE_NAMES.items()}
NEN = {TRI: 3, QUA: 4, TET: 4, HEX: 8, COH: 4}
NIP = {TRI: 1, QUA: 4, TET: 1, HEX: 8, COH: 2}
DIM = {TRI: 2, QUA: 2, TET: 3, HEX: 3, COH: 2}


@dataclasses.dataclass
class Mesh:
    pdim: int
    etype: np.ndarray          # [nels] int32 type codes
    eptr: np.ndarray           # [nels+1] int32
    eind: np.ndarray           # 1-based LOCAL node ids
    emat: np.ndarray           # [nels] 1-based material row, 0 for coh
    enlmat: np.ndarray         # [nels] 0 linear / k -> cohprops row k
    coords: np.ndarray         # [nlnds, pdim]
    nl2g: np.ndarray           # [nlnds] 1-based global node id
    nnds: int                  # global node count
    mat: np.ndarray            # [nmts, 5] E nu visc expn density
    cohprops: np.ndarray       # [ncoh, 6]
    bcnode: np.ndarray         # [nbcs, 1+pdim] global node, flags (0 = constrained)
    bcval: np.ndarray          # [nbcs, pdim]
    fnode: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros(0, np.int32))
    fval: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros((0, 0)))
    telsd: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros((0, 2), np.int32))   # [ntrcs, 2] element (1-based), side
    tval: np.ndarray = dataclasses.field(default_factory=lambda: np.zeros((0, 0)))              # [ntrcs, pdim+2] components, t1, t2
    t: float = 1.0
    dt: float = 1.0
    nodal_bw: int = 10
    stype: str = "implicit"
    eglobal: Optional[np.ndarray] = None   # [nels] 0-based global element id (rank-local meshes)

    @property
    def nels(self) -> int:
        return len(self.etype)

    @property
    def nlnds(self) -> int:
        return self.coords.shape[0]

    @property
    def cpdim(self) -> int:
        return 3 if self.pdim == 2 else 6

    @property
    def ndof(self) -> int:
        return self.pdim * self.nnds

    def elem_nodes(self, e: int) -> np.ndarray:
        return self.eind[self.eptr[e]:self.eptr[e + 1]]

    def normalized(self) -> "Mesh":
        """dtype / contiguity normalisation (what the C ABI expects)."""
        i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
        f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        return dataclasses.replace(
            self, etype=i32(self.etype), eptr=i32(self.eptr), eind=i32(self.eind), emat=i32(self.emat),
            enlmat=i32(self.enlmat), coords=f64(self.coords).reshape(-1, self.pdim), nl2g=i32(self.nl2g),
            mat=f64(self.mat).reshape(-1, 5), cohprops=f64(self.cohprops).reshape(-1, 6),
            bcnode=i32(self.bcnode).reshape(-1, 1 + self.pdim), bcval=f64(self.bcval).reshape(-1, self.pdim))


def from_uniform(pdim, etype_code, conn, coords, E=3.0e10, nu=0.25, **kw) -> Mesh:
    """Single-type mesh from a [nels, nen] 1-based connectivity table."""
    conn = np.asarray(conn, dtype=np.int32)
    nels, nen = conn.shape
    coords = np.asarray(coords, dtype=np.float64).reshape(-1, pdim)
    return Mesh(pdim=pdim, etype=np.full(nels, etype_code, np.int32),
                eptr=np.arange(0, (nels + 1) * nen, nen, dtype=np.int32), eind=conn.reshape(-1).copy(),
                emat=np.ones(nels, np.int32), enlmat=np.zeros(nels, np.int32), coords=coords,
                nl2g=np.arange(1, coords.shape[0] + 1, dtype=np.int32), nnds=coords.shape[0],
                mat=np.array([[E, nu, 1.0e18, 1.0, 3000.0]]), cohprops=np.zeros((0, 6)),
                bcnode=np.zeros((0, 1 + pdim), np.int32), bcval=np.zeros((0, pdim)), **kw).normalized()


# --------------------------------------------------------------------------- .inp grammar
class _Records:
    """Fortran list-directed reads: each read starts on a fresh record and pulls records
    until it has the items it needs; anything left on the last record is dropped (that is
    why trailing `! comments` work in the decks)."""

    def __init__(self, text: str):
        self.lines = text.splitlines()
        self.pos = 0

    def read(self, n: int):
        out = []
        while len(out) < n:
            if self.pos >= len(self.lines):
                raise ValueError("unexpected end of input deck")
            line = self.lines[self.pos].split("!")[0].replace(",", " ")
            self.pos += 1
            out.extend(line.split())
        return out[:n]

    def peek_first(self):
        p = self.pos
        while True:
            toks = self.lines[p].split("!")[0].replace(",", " ").split()
            if toks:
                return toks[0]
            p += 1


def _f(tok: str) -> float:
    return float(tok.replace("d", "e").replace("D", "E"))


def read_inp(path_or_text: str) -> Mesh:
    """Parse a deck in the CURRENT grammar (8 header counts; elements carry `mat nlMat`)."""
    text = path_or_text
    if "\n" not in path_or_text:
        with open(path_or_text) as fh:
            text = fh.read()
    r = _Records(text)
    stype, pdim, nodal_bw = r.read(3)
    pdim, nodal_bw = int(pdim), int(nodal_bw)
    nels, nnds, nmts, ncoh, nceqs, nfrcs, ntrcs, nbcs = (int(x) for x in r.read(8))
    t, dt = (_f(x) for x in r.read(2))
    etype, eptr, eind, emat, enl = [], [0], [], [], []
    for _ in range(nels):
        tag = r.peek_first().strip("'\"")
        code = TYPE_NAMES[tag]
        toks = r.read(1 + NEN[code] + 2)
        etype.append(code)
        eind.extend(int(x) for x in toks[1:1 + NEN[code]])
        eptr.append(len(eind))
        emat.append(int(toks[-2]))
        enl.append(int(toks[-1]))
    coords = np.array([[_f(x) for x in r.read(pdim)] for _ in range(nnds)]).reshape(nnds, pdim)
    mat = np.array([[_f(x) for x in r.read(5)] for _ in range(nmts)]).reshape(nmts, 5)
    coh = np.zeros((ncoh, 6))
    for i in range(ncoh):
        toks = r.read(7)  # seplaw id then PROP_COUNTS(1) = 6 props (src/m_seplaw.F90:8)
        coh[i] = [_f(x) for x in toks[1:]]
    # order in the file follows the reads in main: BCs, forces, tractions (src/main.F90:100-102)
    bcnode = np.zeros((nbcs, 1 + pdim), np.int32)
    bcval = np.zeros((nbcs, pdim))
    for i in range(nbcs):
        toks = r.read(1 + 2 * pdim)
        bcnode[i] = [int(x) for x in toks[:1 + pdim]]
        bcval[i] = [_f(x) for x in toks[1 + pdim:]]
    fnode = np.zeros(nfrcs, np.int32)
    fval = np.zeros((nfrcs, pdim + 2))
    for i in range(nfrcs):
        toks = r.read(1 + pdim + 2)
        fnode[i] = int(toks[0])
        fval[i] = [_f(x) for x in toks[1:]]
    telsd = np.zeros((ntrcs, 2), np.int32)               # ReadDistTractions, src/m_io.F90:436-439
    tval = np.zeros((ntrcs, pdim + 2))
    for i in range(ntrcs):
        toks = r.read(2 + pdim + 2)
        telsd[i] = [int(toks[0]), int(toks[1])]
        tval[i] = [_f(x) for x in toks[2:]]
    return Mesh(pdim=pdim, etype=np.array(etype), eptr=np.array(eptr), eind=np.array(eind),
                emat=np.array(emat), enlmat=np.array(enl), coords=coords,
                nl2g=np.arange(1, nnds + 1), nnds=nnds, mat=mat, cohprops=coh, bcnode=bcnode,
                bcval=bcval, fnode=fnode, fval=fval, telsd=telsd, tval=tval, t=t, dt=dt, nodal_bw=nodal_bw,
                stype=stype.strip("'\"")).normalized()


def write_inp(m: Mesh, path: str) -> None:
    """Emit the current grammar (global mesh: nl2g must be the identity)."""
    assert np.array_equal(m.nl2g, np.arange(1, m.nnds + 1)), "write_inp wants a global mesh"
    with open(path, "w") as fh:
        fh.write(f"{m.stype} {m.pdim} {m.nodal_bw}\n")
        fh.write(f"{m.nels} {m.nnds} {len(m.mat)} {len(m.cohprops)} 0 {len(m.fnode)} {len(m.telsd)} {len(m.bcnode)}\n")
        fh.write(f"{m.t!r} {m.dt!r}\n")
        for e in range(m.nels):
            nodes = " ".join(str(int(x)) for x in m.elem_nodes(e))
            fh.write(f"{TYPE_TAGS[int(m.etype[e])]} {nodes} {int(m.emat[e])} {int(m.enlmat[e])}\n")
        for x in m.coords:
            fh.write(" ".join(repr(float(v)) for v in x) + "\n")
        for row in m.mat:
            fh.write(" ".join(repr(float(v)) for v in row) + "\n")
        for row in m.cohprops:
            fh.write("1 " + " ".join(repr(float(v)) for v in row) + "\n")
        for i in range(len(m.bcnode)):
            fh.write(" ".join(str(int(v)) for v in m.bcnode[i]) + " " +
                     " ".join(repr(float(v)) for v in m.bcval[i]) + "\n")
        for i in range(len(m.fnode)):
            fh.write(f"{int(m.fnode[i])} " + " ".join(repr(float(v)) for v in m.fval[i]) + "\n")
        for i in range(len(m.telsd)):
            fh.write(f"{int(m.telsd[i][0])} {int(m.telsd[i][1])} " + " ".join(repr(float(v)) for v in m.tval[i]) + "\n")


# --------------------------------------------------------------------------- VTK writer
VTK_ID = {TRI: 5, QUA: 9, TET: 10, HEX: 12, COH: 9}     # getVtkid, src/m_elems.F90:108-120


def fortran_f0(v: float, d: int) -> str:
    """Fortran `F0.d` edit descriptor as gfortran prints it: minimal width, no leading zero for |v| < 1
    (`.050000`, `-.000000`), sign kept for negative zero and for negatives that round to zero."""
    t = f"{float(v):.{d}f}"
    if t.startswith("0."):
        return t[1:]
    if t.startswith("-0."):
        return "-" + t[2:]
    return t


def write_vtk(m: Mesh, stress_node: np.ndarray, u_local: np.ndarray, path: Optional[str] = None) -> str:
    """WriteOutput (src/m_io.F90:480-555): legacy ASCII VTK of one rank -- local points (F0.3, z = 0 appended in
    2-D), cells with 0-based local node ids, nodal STRESS (cpdim components, F0.6) and displacement vectors
    (F0.6).  Byte-for-byte what the reference writes (tests/test_host_cpu.py checks the golden files); compare
    ARRAYS for the 1e-10 field tolerance, the text carries six decimals.  Returns the text; writes it if `path`."""
    pd, n = m.pdim, m.nlnds
    s = np.asarray(stress_node, float).reshape(n, m.cpdim)
    u = np.asarray(u_local, float).reshape(n, pd)
    pad = [0.0] if pd == 2 else []
    out = ["# vtk DataFile Version 2.0", "File written by Defmod", "ASCII", "DATASET UNSTRUCTURED_GRID", f"POINTS {n} double"]
    out += [" ".join(fortran_f0(v, 3) for v in list(x) + pad) for x in m.coords]
    out.append(f"CELLS {m.nels} {int(m.eptr[-1]) + m.nels}")
    for e in range(m.nels):
        nd = m.elem_nodes(e)
        out.append(" ".join(str(int(v)) for v in [len(nd)] + list(nd - 1)))
    out.append(f"CELL_TYPES {m.nels}")
    out += [str(VTK_ID[int(t)]) for t in m.etype]
    out += [f"POINT_DATA {n}", f"SCALARS STRESS FLOAT {m.cpdim}", "LOOKUP_TABLE DEFAULT"]
    out += [" ".join(fortran_f0(v, 6) for v in row) for row in s]
    out.append("VECTORS displacements double")
    out += [" ".join(fortran_f0(v, 6) for v in list(row) + pad) for row in u]
    text = "\n".join(out) + "\n"
    if path:
        with open(path, "w") as fh:
            fh.write(text)
    return text


def read_vtk(text: str):
    """Parse a WriteOutput file back: (coords[n, 3], cells list of 0-based node lists, cell types, stress[n, cp], disp[n, 3])."""
    tok = text.split("\n")
    i = next(k for k, l in enumerate(tok) if l.startswith("POINTS"))
    n = int(tok[i].split()[1])
    pts = np.array([[float(v) for v in tok[i + 1 + k].split()] for k in range(n)])
    i = next(k for k, l in enumerate(tok) if l.startswith("CELLS"))
    ne = int(tok[i].split()[1])
    cells = [[int(v) for v in tok[i + 1 + k].split()[1:]] for k in range(ne)]
    i = next(k for k, l in enumerate(tok) if l.startswith("CELL_TYPES"))
    types = [int(tok[i + 1 + k]) for k in range(ne)]
    i = next(k for k, l in enumerate(tok) if l.startswith("LOOKUP_TABLE"))
    stress = np.array([[float(v) for v in tok[i + 1 + k].split()] for k in range(n)])
    i = next(k for k, l in enumerate(tok) if l.startswith("VECTORS"))
    disp = np.array([[float(v) for v in tok[i + 1 + k].split()] for k in range(n)])
    return pts, cells, types, stress, disp


# --------------------------------------------------------------------------- distribution
def localize(g: Mesh, epart: np.ndarray, rank: int) -> Mesh:
    """Rank-local mesh exactly as src/m_io.F90:200-237 + src/main.F90:49-96 build it:
    elements with epart == rank in ascending global order; local node ids assigned in
    ascending global node order; nl2g the inverse map."""
    epart = np.asarray(epart)
    sel = np.nonzero(epart == rank)[0]
    counts = (g.eptr[1:] - g.eptr[:-1])[sel]
    eptr = np.zeros(len(sel) + 1, np.int64)
    np.cumsum(counts, out=eptr[1:])
    idx = np.repeat(g.eptr[sel].astype(np.int64) - eptr[:-1], counts) + np.arange(eptr[-1])
    gnodes = g.eind[idx]
    nl2g = np.unique(gnodes)                      # ascending global ids -> local 1..nlnds
    work = np.zeros(g.nnds + 1, np.int32)
    work[nl2g] = np.arange(1, len(nl2g) + 1)
    fnode, fval, telsd, tval = distribute_loads(g, epart, int(epart.max()) + 1, rank)
    return dataclasses.replace(
        g, etype=g.etype[sel], eptr=eptr.astype(np.int32), eind=work[gnodes], emat=g.emat[sel],
        enlmat=g.enlmat[sel], coords=g.coords[nl2g - 1], nl2g=nl2g.astype(np.int32),
        eglobal=sel.astype(np.int64), fnode=fnode, fval=fval, telsd=telsd, tval=tval).normalized()


def distribute_loads(g: Mesh, epart: np.ndarray, nprcs: int, rank: int):
    """What ReadDistForces / ReadDistTractions leave on `rank` (src/m_io.F90:330-372, :414-475): nodal forces dealt
    round-robin (item i goes to rank i mod nprcs, GLOBAL node ids kept), tractions follow their element's partition
    with the element id renumbered to the rank-local one (emap).  Returns fnode, fval, telsd, tval for femx_set_loads."""
    epart = np.asarray(epart)
    fnode, fval = g.fnode[rank::nprcs], np.asarray(g.fval)[rank::nprcs]
    telsd, tval = np.asarray(g.telsd).reshape(-1, 2), np.asarray(g.tval)
    emap = np.zeros(len(epart) + 1, np.int32)
    mine = np.flatnonzero(epart == rank)
    emap[mine + 1] = np.arange(1, len(mine) + 1)
    keep = epart[telsd[:, 0] - 1] == rank if len(telsd) else np.zeros(0, bool)
    loc = telsd[keep].copy()
    if len(loc):
        loc[:, 0] = emap[loc[:, 0]]
    return fnode, fval, loc, tval[keep] if len(telsd) else tval


def renumber_by_partition(g: Mesh, epart: np.ndarray, nparts: int):
    """Global node numbering that follows an element partition (what a production host does before it creates its PETSc
    objects; the reference keeps the mesh file's numbering, src/main.F90:57-63, so with a METIS partition most rows a
    rank assembles belong to other ranks).  A node goes to the lowest part among the elements that touch it; nodes are
    numbered part by part, old order inside a part.  Returns (mesh with permuted node ids, node_bounds[nparts + 1]):
    part q holds the new nodes node_bounds[q] + 1 .. node_bounds[q + 1], i.e. the rows pdim * node_bounds[q] ..
    pdim * node_bounds[q + 1] - 1 -- pass pdim * node_bounds as `row_bounds` to parallel.DistributedAssembly (PETSc:
    the local sizes of MatCreateAIJ / VecCreateMPI instead of PETSC_DECIDE)."""
    epart = np.asarray(epart, np.int64)
    npart = np.full(g.nnds + 1, nparts, np.int64)
    counts = (g.eptr[1:] - g.eptr[:-1]).astype(np.int64)
    np.minimum.at(npart, np.asarray(g.eind, np.int64), np.repeat(epart, counts))
    npart = npart[1:]
    npart[npart == nparts] = nparts - 1                     # nodes no element touches go last
    order = np.argsort(npart, kind="stable")                # new position -> old node (0-based)
    new_of_old = np.empty(g.nnds, np.int64)
    new_of_old[order] = np.arange(g.nnds)
    bounds = np.searchsorted(npart[order], np.arange(nparts + 1)).astype(np.int64)
    relabel = lambda ids: (new_of_old[np.asarray(ids, np.int64) - 1] + 1).astype(np.int32)   # noqa: E731
    bcnode = np.array(g.bcnode, np.int32).copy()
    if len(bcnode):
        bcnode[:, 0] = relabel(bcnode[:, 0])
        o = np.argsort(bcnode[:, 0], kind="stable")
        bcnode, bcval = bcnode[o], np.asarray(g.bcval)[o]
    else:
        bcval = g.bcval
    assert np.array_equal(g.nl2g, np.arange(1, g.nnds + 1)), "renumber the global mesh, before localize"
    g2 = dataclasses.replace(g, eind=relabel(g.eind), coords=np.asarray(g.coords)[order], bcnode=bcnode, bcval=bcval,
                             fnode=relabel(g.fnode) if len(g.fnode) else g.fnode).normalized()
    return g2, bounds


def petsc_ownership(n: int, nranks: int, rank: int):
    """PETSC_DECIDE row split used by MatCreateAIJ/VecCreateMPI (src/main.F90:137-138,154):
    rank r owns n/P + (r < n mod P) consecutive rows.  Returns [start, end)."""
    base, rem = divmod(n, nranks)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def metis_partition(g: Mesh, nparts: int):
    """Element partition the way PartitionBroadcast gets it (src/m_io.F90:125-141): METIS_PartMeshNodal on the 0-based
    element -> node lists, every optional argument null, epart kept.  Calls tools/libmetispart.so (tools/metis_part.c: the
    METIS inside the CUDA toolkit, 64-bit idx_t), built on first use with `make -C tools metis`.  Returns (epart, edge cut);
    raises RuntimeError when the wrapper cannot be built or METIS fails."""
    import ctypes
    import os
    import subprocess
    tools = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools")
    so = os.path.join(tools, "libmetispart.so")
    if not os.path.exists(so):
        r = subprocess.run(["make", "-C", tools, "metis"], capture_output=True, text=True)
        if r.returncode != 0 or not os.path.exists(so):
            raise RuntimeError("cannot build tools/libmetispart.so: " + (r.stderr or r.stdout)[-500:])
    lib = ctypes.CDLL(so)
    fn = lib.femx_metis_part_mesh_nodal
    ip = ctypes.POINTER(ctypes.c_int)
    fn.argtypes = [ctypes.c_int, ctypes.c_int, ip, ip, ctypes.c_int, ip, ctypes.POINTER(ctypes.c_longlong)]
    fn.restype = ctypes.c_int
    eptr = np.ascontiguousarray(g.eptr, np.int32)
    eind = np.ascontiguousarray(np.asarray(g.eind, np.int32) - 1)
    epart = np.zeros(g.nels, np.int32)
    cut = ctypes.c_longlong(0)
    rc = fn(g.nels, g.nnds, eptr.ctypes.data_as(ip), eind.ctypes.data_as(ip), int(nparts), epart.ctypes.data_as(ip), ctypes.byref(cut))
    if rc != 1:
        raise RuntimeError(f"METIS_PartMeshNodal returned {rc}")
    return epart, int(cut.value)
