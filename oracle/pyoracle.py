"""ctypes view of oracle/liboracle.so (the CPU restatement; see oracle.h).

TEST INFRASTRUCTURE ONLY.  Parity pinned by tests/test_oracle_golden.py against the
reference's fixtures (examples/tmp.txt, elastic golden VTK, UEL traction printout).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

TRI, QUA, TET, HEX, COH = 1, 2, 3, 4, 5
Q_GATHER_LOCAL14, Q_TANGENT_PRECEDENCE, Q_FORCE_LAST_IP, Q_COH_COUNTS_NODE = 1, 2, 4, 8
Q_REFERENCE, Q_UEL = 15, 0

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


class _Mesh(C.Structure):
    _fields_ = [("pdim", C.c_int), ("nels", C.c_int), ("etype", _ip), ("eptr", _ip), ("eind", _ip),
                ("emat", _ip), ("enlmat", _ip), ("nlnds", C.c_int), ("coords", _dp), ("nl2g", _ip),
                ("nnds", C.c_int), ("nmts", C.c_int), ("mat", _dp), ("ncoh", C.c_int),
                ("cohprops", _dp), ("nbcs", C.c_int), ("bcnode", _ip), ("bcval", _dp)]


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = [os.path.join(_HERE, f) for f in ("oracle.cpp", "oracle.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.orc_csr_create.restype = C.c_void_p
        L.orc_csr_copy.restype = C.c_void_p
        L.orc_assemble_K.restype = C.c_void_p
        L.orc_jacobian.restype = C.c_void_p
        L.orc_csr_nnz.restype = C.c_int64
        L.orc_time_assembly.restype = C.c_double
        _LIB = L
    return _LIB


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def nen(t):
    return lib().orc_nen(t)


def nip(t):
    return lib().orc_nip(t)


def dim(t):
    return lib().orc_dim(t)


def sampts(t):
    pts = np.zeros((nip(t), dim(t)))
    w = np.zeros(nip(t))
    lib().orc_sampts(t, _d(pts), _d(w))
    return pts, w


def shape(t, ip):
    N = np.zeros(nen(t))
    lib().orc_shape(t, ip, _d(N))
    return N


def shape_d(t, pt):
    dN = np.zeros((dim(t), nen(t)))
    pt = np.ascontiguousarray(pt, dtype=np.float64)
    lib().orc_shape_d(t, _d(pt), _d(dN))
    return dN


def nodal_stress_inv(t):
    inv = np.zeros((nen(t), nip(t)))
    ok = lib().orc_nodal_stress_inv(t, _d(inv))
    return inv if ok else None


def form_el_ke(t, ecoords, E, nu):
    """FormElKE; returns Ke as a (ndof, ndof) array with Ke[i, j] = k(i+1, j+1)."""
    ecoords = np.ascontiguousarray(ecoords, dtype=np.float64)
    nd = nen(t) * dim(t)
    k = np.zeros((nd, nd))
    lib().orc_form_el_ke(t, _d(ecoords), C.c_double(E), C.c_double(nu), _d(k))
    return k.T.copy()  # column-major memory -> [i, j]


def calc_el_stress(t, ecoords, edisp, E, nu):
    ecoords = np.ascontiguousarray(ecoords, dtype=np.float64)
    edisp = np.ascontiguousarray(edisp, dtype=np.float64)
    cp = 3 if dim(t) == 2 else 6
    s = np.zeros((nip(t), cp))
    lib().orc_calc_el_stress(t, _d(ecoords), _d(edisp), C.c_double(E), C.c_double(nu), _d(s))
    return s


def recover_nodal_stress(t, estress):
    estress = np.ascontiguousarray(estress, dtype=np.float64)
    cp = 3 if dim(t) == 2 else 6
    out = np.zeros((nen(t), cp))
    lib().orc_recover_nodal_stress(t, _d(estress), _d(out))
    return out


def seplaw_1_tract(props, gap, vgap, dt):
    props, gap, vgap = (np.ascontiguousarray(x, dtype=np.float64) for x in (props, gap, vgap))
    out = np.zeros(2)
    lib().orc_seplaw_1_tract(_d(props), _d(gap), _d(vgap), C.c_double(dt), _d(out))
    return out


def seplaw_1_stiff(props, gap, vgap, dt):
    props, gap, vgap = (np.ascontiguousarray(x, dtype=np.float64) for x in (props, gap, vgap))
    out = np.zeros((2, 2))
    lib().orc_seplaw_1_stiff(_d(props), _d(gap), _d(vgap), C.c_double(dt), _d(out))
    return out


def apply_tract_1(ecoords, nodes, u_local, dt, props, quirks, per_ip=False):
    ecoords = np.ascontiguousarray(ecoords, dtype=np.float64)
    nodes = np.ascontiguousarray(nodes, dtype=np.int32)
    u_local = np.ascontiguousarray(u_local, dtype=np.float64)
    props = np.ascontiguousarray(props, dtype=np.float64)
    out = np.zeros(8)
    pip = np.zeros((2, 8))
    lib().orc_apply_tract_1(_d(ecoords), _i(nodes), _d(u_local), C.c_double(dt), _d(props),
                            int(quirks), _d(out), _d(pip))
    return (out, pip) if per_ip else out


def apply_stiff_1(ecoords, nodes, u_local, dt, props, quirks):
    ecoords = np.ascontiguousarray(ecoords, dtype=np.float64)
    nodes = np.ascontiguousarray(nodes, dtype=np.int32)
    u_local = np.ascontiguousarray(u_local, dtype=np.float64)
    props = np.ascontiguousarray(props, dtype=np.float64)
    out = np.zeros((8, 8))
    lib().orc_apply_stiff_1(_d(ecoords), _i(nodes), _d(u_local), C.c_double(dt), _d(props),
                            int(quirks), _d(out))
    return out


class Mesh:
    """Keeps the numpy arrays alive behind an orc_mesh struct.

    Accepts any object with the fem_b200.mesh.Mesh attribute names (duck-typed so the
    oracle does not import the product package).
    """

    def __init__(self, m):
        self.pdim = int(m.pdim)
        self.etype = np.ascontiguousarray(m.etype, dtype=np.int32)
        self.eptr = np.ascontiguousarray(m.eptr, dtype=np.int32)
        self.eind = np.ascontiguousarray(m.eind, dtype=np.int32)
        self.emat = np.ascontiguousarray(m.emat, dtype=np.int32)
        self.enlmat = np.ascontiguousarray(m.enlmat, dtype=np.int32)
        self.coords = np.ascontiguousarray(m.coords, dtype=np.float64)
        self.nl2g = np.ascontiguousarray(m.nl2g, dtype=np.int32)
        self.nnds = int(m.nnds)
        self.mat = np.ascontiguousarray(np.asarray(m.mat, dtype=np.float64).reshape(-1, 5)[:, :2]) \
            if np.asarray(m.mat).size and np.asarray(m.mat).reshape(len(m.mat), -1).shape[1] == 5 \
            else np.ascontiguousarray(m.mat, dtype=np.float64).reshape(-1, 2)
        self.cohprops = np.ascontiguousarray(m.cohprops, dtype=np.float64).reshape(-1, 6)
        self.bcnode = np.ascontiguousarray(m.bcnode, dtype=np.int32).reshape(-1, 1 + self.pdim)
        self.bcval = np.ascontiguousarray(m.bcval, dtype=np.float64).reshape(-1, self.pdim)
        self.nels = len(self.etype)
        self.nlnds = self.coords.shape[0]
        self.c = _Mesh(self.pdim, self.nels, _i(self.etype), _i(self.eptr), _i(self.eind),
                       _i(self.emat), _i(self.enlmat), self.nlnds, _d(self.coords), _i(self.nl2g),
                       self.nnds, len(self.mat), _d(self.mat), len(self.cohprops), _d(self.cohprops),
                       len(self.bcnode), _i(self.bcnode), _d(self.bcval))

    @property
    def ndof(self):
        return self.pdim * self.nnds


class Csr:
    def __init__(self, handle):
        self.h = C.c_void_p(handle)

    def __del__(self):
        if self.h:
            lib().orc_csr_destroy(self.h)
            self.h = None

    def export(self, nrows):
        nnz = lib().orc_csr_nnz(self.h)
        rowptr = np.zeros(nrows + 1, dtype=np.int64)
        colidx = np.zeros(nnz, dtype=np.int32)
        vals = np.zeros(nnz)
        lib().orc_csr_export(self.h, rowptr.ctypes.data_as(C.POINTER(C.c_int64)),
                             colidx.ctypes.data_as(C.POINTER(C.c_int32)), _d(vals))
        return rowptr, colidx, vals


def assemble_K(mesh, prealloc=0, apply_bc=False):
    return Csr(lib().orc_assemble_K(C.byref(mesh.c), int(prealloc), int(apply_bc)))


def jacobian(mesh, K, u_total_local, dt, quirks):
    u = np.ascontiguousarray(u_total_local, dtype=np.float64)
    return Csr(lib().orc_jacobian(C.byref(mesh.c), K.h, _d(u), C.c_double(dt), int(quirks)))


def residual(mesh, J, u_total_local, du_global, f_ext_global, dt, t_total, quirks):
    u = np.ascontiguousarray(u_total_local, dtype=np.float64)
    du = np.ascontiguousarray(du_global, dtype=np.float64)
    f = np.ascontiguousarray(f_ext_global, dtype=np.float64)
    r = np.zeros(mesh.ndof)
    lib().orc_residual(C.byref(mesh.c), J.h, _d(u), _d(du), _d(f), C.c_double(dt),
                       C.c_double(t_total), int(quirks), _d(r))
    return r


def form_rhs(mesh, fnode, fval, telsd, tval, t_init, dt, t_total):
    """FormRHS (src/m_global.F90:373-436) on the rank-local mesh: global-length Vec_F."""
    fn = np.ascontiguousarray(fnode, dtype=np.int32).reshape(-1)
    fv = np.ascontiguousarray(fval, dtype=np.float64).reshape(len(fn), -1) if len(fn) else np.zeros((0, mesh.pdim + 2))
    te = np.ascontiguousarray(telsd, dtype=np.int32).reshape(-1, 2)
    tv = np.ascontiguousarray(tval, dtype=np.float64).reshape(len(te), -1) if len(te) else np.zeros((0, mesh.pdim + 2))
    f = np.zeros(mesh.ndof)
    lib().orc_form_rhs(C.byref(mesh.c), len(fn), _i(fn), _d(fv), len(te), _i(te), _d(tv), C.c_double(t_init), C.c_double(dt),
                       C.c_double(t_total), _d(f))
    return f


def edge_area_nodes(etype, ecoords, side):
    x = np.ascontiguousarray(ecoords, dtype=np.float64)
    area = C.c_double()
    slots = (C.c_int * 4)()
    nps = lib().orc_edge_area_nodes(int(etype), _d(x), int(side), C.byref(area), slots)
    return area.value, list(slots[:nps])


def fint(mesh, u_local):
    u = np.ascontiguousarray(u_local, dtype=np.float64)
    f = np.zeros(mesh.ndof)
    lib().orc_fint(C.byref(mesh.c), _d(u), _d(f))
    return f


def stress(mesh, uu_local, quirks=Q_REFERENCE):
    u = np.ascontiguousarray(uu_local, dtype=np.float64)
    cp = 3 if mesh.pdim == 2 else 6
    s = np.zeros((mesh.nlnds, cp))
    cnt = np.zeros(mesh.nlnds, dtype=np.int32)
    lib().orc_stress(C.byref(mesh.c), _d(u), int(quirks), _d(s), _i(cnt))
    return s, cnt


def time_assembly(mesh, u_local, prealloc=0, nthreads=1):
    u = np.ascontiguousarray(u_local, dtype=np.float64)
    nnz = C.c_int64(0)
    sec = lib().orc_time_assembly(C.byref(mesh.c), int(prealloc), _d(u), int(nthreads), C.byref(nnz))
    return sec, nnz.value
