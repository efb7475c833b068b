"""Host-side mirror of the reference's calling surface over the femx C ABI.

The reference is Fortran; no Fortran compiler exists in this image, so the host layer a
`defmod` maintainer would write in Fortran (fortran/femx_mod.F90) is
mirrored here in Python with the same procedure names and argument meaning:

  FormElKE      src/m_local.F90:21      CalcElStress        src/m_local.F90:93
  FormLocalK    src/m_global.F90:239    RecoverNodalStress  src/m_global.F90:488
  applyTract_1  src/m_global.F90:622    applyStiff_1        src/m_global.F90:762
  CalcJacobian  src/m_global.F90:98     CalcResidual        src/m_global.F90:161
  the K loop    src/main.F90:157-171    the stress loops    src/main.F90:253-279

Every call goes through libfemx.so; nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C
import dataclasses

import numpy as np

from . import _lib
from .mesh import DIM, NEN, NIP, Mesh

HOST, DEVICE = 0, 1
ROW_MAJOR, COL_MAJOR = 0, 1
Q_GATHER_LOCAL14, Q_TANGENT_PRECEDENCE, Q_FORCE_LAST_IP, Q_COH_COUNTS_NODE = 1, 2, 4, 8
Q_REFERENCE, Q_UEL = 15, 0


class FemxError(RuntimeError):
    pass


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _dp(a):
    return a.ctypes.data_as(_lib.c_double_p)


def _ip(a):
    return a.ctypes.data_as(_lib.c_int_p)


class DeviceBuffer:
    """Raw device allocation owned by a Femx context (for device-resident calls)."""

    def __init__(self, fx: "Femx", nbytes: int):
        self.fx, self.nbytes = fx, int(nbytes)
        ptr = C.c_void_p()
        fx._ck(fx.L.femx_device_alloc(fx.h, C.byref(ptr), self.nbytes))
        self.ptr = ptr

    def upload(self, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        assert arr.nbytes <= self.nbytes
        self.fx._ck(self.fx.L.femx_memcpy(self.fx.h, self.ptr, _p(arr), arr.nbytes, DEVICE, HOST))
        return self

    def download(self, dtype=np.float64, count=None) -> np.ndarray:
        n = self.nbytes // np.dtype(dtype).itemsize if count is None else count
        out = np.empty(n, dtype=dtype)
        self.fx._ck(self.fx.L.femx_memcpy(self.fx.h, _p(out), self.ptr, out.nbytes, HOST, DEVICE))
        return out

    def free(self):
        if self.ptr:
            self.fx.L.femx_device_free(self.fx.h, self.ptr)
            self.ptr = None


class PinnedArray:
    """numpy view of pinned host memory from femx_host_alloc."""

    def __init__(self, L, n: int, dtype=np.float64):
        self.L = L
        self.ptr = C.c_void_p()
        nbytes = int(n) * np.dtype(dtype).itemsize
        if L.femx_host_alloc(C.byref(self.ptr), nbytes) != 0:
            raise FemxError("femx_host_alloc failed")
        buf = (C.c_char * max(nbytes, 1)).from_address(self.ptr.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(n))

    def free(self):
        if self.ptr:
            self.array = None
            self.L.femx_host_free(self.ptr)
            self.ptr = None


class Femx:
    def __init__(self, device: int = 0, quirks: int = Q_REFERENCE, assembly_variant: int = 0, tile: int = 0, lz: int = 0, dbg: int = 0,
                 peer_in_kernel: bool = False, host_plan: bool = False):
        self.L = _lib.lib()
        self.h = C.c_void_p()
        opts = _lib.FemxOpts(quirks=quirks, assembly_variant=assembly_variant)
        opts.reserved[0], opts.reserved[1], opts.reserved[2] = tile, lz, dbg     # fused-kernel tile shape / z-chunk (tuning knobs)
        opts.reserved[3] = 1 if peer_in_kernel else 0   # peer exchange: the assembly kernel stores interface rows at their owner itself
        opts.reserved[4] = 1 if host_plan else 0        # integer pass on the host (plan.cpp: the cross-check of the device pass)
        rc = self.L.femx_create(C.byref(self.h), device, C.byref(opts))
        if rc != 0:
            raise FemxError(f"femx_create failed ({rc}): {self.L.femx_last_error(None).decode()}")
        self.mesh = None
        self.nrows = self.nnz = 0

    # ------------------------------------------------------------------ plumbing
    def _ck(self, rc):
        if rc != 0:
            raise FemxError(f"femx error {rc}: {self.L.femx_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            self.L.femx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def launches(self) -> int:
        return self.L.femx_launch_count(self.h)

    @property
    def last_kernel_ms(self) -> float:
        return self.L.femx_last_kernel_ms(self.h)

    def device_buffer(self, nbytes) -> DeviceBuffer:
        return DeviceBuffer(self, nbytes)

    def pinned(self, n, dtype=np.float64) -> PinnedArray:
        return PinnedArray(self.L, n, dtype)

    # ------------------------------------------------------------------ state
    def set_mesh(self, m: Mesh, coords_layout=ROW_MAJOR):
        m = m.normalized()
        self.mesh = m
        coords = m.coords if coords_layout == ROW_MAJOR else np.asfortranarray(m.coords)
        self._keep = coords
        self._ck(self.L.femx_set_mesh(self.h, m.pdim, m.nels, _ip(m.etype), _ip(m.eptr), _ip(m.eind), _ip(m.emat),
                                      _ip(m.enlmat), m.nlnds, coords.ctypes.data_as(_lib.c_double_p), coords_layout,
                                      _ip(m.nl2g), m.nnds))
        enu = np.ascontiguousarray(m.mat[:, :2])
        coh = np.ascontiguousarray(m.cohprops)
        self._ck(self.L.femx_set_materials(self.h, len(enu), _dp(enu), len(coh), _dp(coh)))
        self._ck(self.L.femx_set_bcs(self.h, len(m.bcnode), _ip(m.bcnode), _dp(m.bcval)))
        return self

    def set_materials(self, mat: np.ndarray, cohprops: np.ndarray | None = None):
        """mat(:, 1:2) = (E, nu) rows (src/m_global.F90:254-255); may be called again after build_plan."""
        enu = np.ascontiguousarray(np.asarray(mat, float)[:, :2])
        coh = np.ascontiguousarray(self.mesh.cohprops if cohprops is None else cohprops)
        self._ck(self.L.femx_set_materials(self.h, len(enu), _dp(enu), len(coh), _dp(coh)))
        self.mesh = dataclasses.replace(self.mesh, mat=np.asarray(mat, float), cohprops=coh)
        return self

    def build_plan(self):
        self._ck(self.L.femx_build_plan(self.h))
        nr, nz = C.c_int64(), C.c_int64()
        self._ck(self.L.femx_pattern_size(self.h, C.byref(nr), C.byref(nz)))
        self.nrows, self.nnz = nr.value, nz.value
        return self

    def pattern(self):
        rowptr = np.empty(self.nrows + 1, np.int64)
        colidx = np.empty(self.nnz, np.int32)
        self._ck(self.L.femx_get_pattern(self.h, _p(rowptr), _p(colidx), HOST))
        return rowptr, colidx

    def slot_map(self):
        off = np.empty(self.mesh.nels + 1, np.int64)
        self._ck(self.L.femx_get_slot_map(self.h, off.ctypes.data_as(_lib.c_i64_p), None))
        bpos = np.empty(off[-1], np.uint16)
        self._ck(self.L.femx_get_slot_map(self.h, off.ctypes.data_as(_lib.c_i64_p), bpos.ctypes.data_as(_lib.c_u16_p)))
        return off, bpos

    # ------------------------------------------------------------------ element level
    def FormElKE(self, eltype: int, ecoords, E, nu) -> np.ndarray:
        """Batched FormElKE: ecoords [n, nen, dim] (or [nen, dim]); returns k [n, ndof, ndof]
        with k[e, i, j] = Fortran k(i+1, j+1)."""
        x = np.ascontiguousarray(ecoords, dtype=np.float64)
        single = x.ndim == 2
        x = x.reshape(-1, NEN[eltype], DIM[eltype])
        n = len(x)
        enu = np.empty((n, 2))
        enu[:, 0], enu[:, 1] = E, nu
        nd = NEN[eltype] * DIM[eltype]
        out = np.empty((n, nd, nd))
        self._ck(self.L.femx_element_ke(self.h, eltype, n, _dp(x), _dp(enu), _dp(out)))
        out = out.transpose(0, 2, 1)  # column-major memory
        return out[0] if single else out

    def CalcElStress(self, eltype: int, ecoords, edisp, E, nu):
        """Batched CalcElStress + RecoverNodalStress: returns (estress [n,nip,cp], nodal [n,nen,cp])."""
        x = np.ascontiguousarray(ecoords, dtype=np.float64)
        single = x.ndim == 2
        x = x.reshape(-1, NEN[eltype], DIM[eltype])
        u = np.ascontiguousarray(edisp, dtype=np.float64).reshape(x.shape)
        n = len(x)
        enu = np.empty((n, 2))
        enu[:, 0], enu[:, 1] = E, nu
        cp = 3 if DIM[eltype] == 2 else 6
        s = np.empty((n, NIP[eltype], cp))
        sn = np.empty((n, NEN[eltype], cp))
        self._ck(self.L.femx_element_stress(self.h, eltype, n, _dp(x), _dp(u), _dp(enu), _dp(s), _dp(sn)))
        return (s[0], sn[0]) if single else (s, sn)

    def cohesive(self, ecoords, nodes, u_local, props, dt):
        """Batched applyTract_1 / applyStiff_1: returns (force [n,8], stiff [n,8,8])."""
        x = np.ascontiguousarray(ecoords, dtype=np.float64).reshape(-1, 4, 2)
        n = len(x)
        nd = np.ascontiguousarray(nodes, dtype=np.int32).reshape(n, 4)
        u = np.ascontiguousarray(u_local, dtype=np.float64).reshape(-1)
        p = np.ascontiguousarray(np.broadcast_to(np.asarray(props, dtype=np.float64), (n, 6)))
        f = np.empty((n, 8))
        s = np.empty((n, 8, 8))
        self._ck(self.L.femx_element_cohesive(self.h, n, _dp(x), _ip(nd), _dp(u), len(u), _dp(p), float(dt), _dp(f), _dp(s)))
        return f, s

    def applyTract_1(self, ecoords, nodes, u_local, dt, props):
        return self.cohesive(ecoords, nodes, u_local, props, dt)[0][0]

    def applyStiff_1(self, ecoords, nodes, u_local, dt, props):
        return self.cohesive(ecoords, nodes, u_local, props, dt)[1][0]

    def FormLocalK(self, el: int):
        """FormLocalK for local element `el` (0-based): (k, indx) with indx the 0-based global
        dofs after `indx=indxmap(indx,2)` (src/main.F90:161)."""
        m = self.mesh
        t = int(m.etype[el])
        nodes = m.elem_nodes(el)
        if m.emat[el] == 0:
            k = np.zeros((NEN[t] * m.pdim,) * 2)
        else:
            E, nu = m.mat[m.emat[el] - 1, :2]
            k = self.FormElKE(t, m.coords[nodes - 1], E, nu)
        indx = (m.pdim * (m.nl2g[nodes - 1].astype(np.int64) - 1)[:, None] + np.arange(m.pdim)).reshape(-1)
        return k, indx

    # ------------------------------------------------------------------ mesh level (host buffers)
    def assemble_K(self) -> np.ndarray:
        vals = np.empty(self.nnz)
        self._ck(self.L.femx_assemble_K(self.h, _p(vals), HOST))
        return vals

    def assemble_K_fint(self, u_local):
        u = np.ascontiguousarray(u_local, dtype=np.float64).reshape(-1)
        vals = np.empty(self.nnz)
        fint = np.empty(self.nrows)
        self._ck(self.L.femx_assemble_K_fint(self.h, _p(u), _p(vals), _p(fint), HOST))
        return vals, fint

    def apply_kbc(self, vals: np.ndarray) -> np.ndarray:
        self._ck(self.L.femx_apply_kbc(self.h, _p(vals), HOST))
        return vals

    def CalcJacobian(self, kvals, u_total_local, dt) -> np.ndarray:
        u = np.ascontiguousarray(u_total_local, dtype=np.float64).reshape(-1)
        k = np.ascontiguousarray(kvals, dtype=np.float64)
        j = np.empty(self.nnz)
        self._ck(self.L.femx_jacobian(self.h, _p(k), _p(u), float(dt), _p(j), HOST))
        return j

    def CalcResidual(self, jvals, u_total_local, du, f_ext, dt, t_total) -> np.ndarray:
        u = np.ascontiguousarray(u_total_local, dtype=np.float64).reshape(-1)
        j = np.ascontiguousarray(jvals, dtype=np.float64)
        du = np.ascontiguousarray(du, dtype=np.float64)
        f = np.ascontiguousarray(f_ext, dtype=np.float64)
        r = np.empty(self.nrows)
        self._ck(self.L.femx_residual(self.h, _p(j), _p(u), _p(du), _p(f), float(dt), float(t_total), _p(r), HOST))
        return r

    def residual_forces(self, u_total_local, f_ext, dt, t_total) -> np.ndarray:
        u = np.ascontiguousarray(u_total_local, dtype=np.float64).reshape(-1)
        f = np.ascontiguousarray(f_ext, dtype=np.float64)
        r = np.empty(self.nrows)
        self._ck(self.L.femx_residual_forces(self.h, _p(u), _p(f), float(dt), float(t_total), _p(r), HOST))
        return r

    def set_loads(self, fnode=(), fval=(), telsd=(), tval=()):
        """fnode/fval, telsd/tval as ReadDistForces / ReadDistTractions leave them (src/m_io.F90:330-372, :414-475)."""
        pd = self.mesh.pdim
        fn = np.ascontiguousarray(fnode, np.int32).reshape(-1)
        fv = np.ascontiguousarray(fval, np.float64).reshape(len(fn), pd + 2) if len(fn) else np.zeros((0, pd + 2))
        te = np.ascontiguousarray(telsd, np.int32).reshape(-1, 2)
        tv = np.ascontiguousarray(tval, np.float64).reshape(len(te), pd + 2) if len(te) else np.zeros((0, pd + 2))
        self._ck(self.L.femx_set_loads(self.h, len(fn), _ip(fn), _dp(fv), len(te), _ip(te), _dp(tv)))
        return self

    def FormRHS(self, t_init, dt, t_total) -> np.ndarray:
        """FormRHS (src/m_global.F90:373-436): global-length Vec_F of this rank."""
        f = np.empty(self.nrows)
        self._ck(self.L.femx_form_rhs(self.h, float(t_init), float(dt), float(t_total), _p(f), HOST))
        return f

    def stress(self, uu_local, layout=ROW_MAJOR) -> np.ndarray:
        """RecoverStress + RecoverNodalStress + average: stress_node [nlnds, cp]."""
        u = np.ascontiguousarray(uu_local, dtype=np.float64).reshape(-1)
        cp = self.mesh.cpdim
        out = np.empty(self.mesh.nlnds * cp)
        self._ck(self.L.femx_stress(self.h, _p(u), _p(out), layout, HOST))
        return out.reshape(cp, -1).T if layout == COL_MAJOR else out.reshape(-1, cp)
