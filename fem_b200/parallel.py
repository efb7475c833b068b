"""Multi-GPU assembly: one rank per GPU, element partition, interface-row exchange.

Mirrors what the reference gets from PETSc for free: every rank assembles the
contributions of ITS elements (METIS `epart`, src/m_io.F90:137-141; here any `epart`), while
matrix rows are owned by the contiguous PETSC_DECIDE split of `MatCreateAIJ`
(src/main.F90:154).  Contributions to rows another rank owns are what
`MatAssemblyBegin/End` ships through its stash (src/main.F90:167-168).

B200 design: the rank-local CSR lists rows in ascending global order, so
  * the values of this rank's OWNED rows are one contiguous slice of the local value array,
  * the values destined to peer q are one contiguous slice as well (no pack kernel: the
    assembly kernel has already written the send buffer),
and the only per-assembly traffic is those fp64 slices, sent with NCCL (`all_to_all_single`
over NVLink; gloo on CPU for the tests).  Integer (row, col) lists travel once, at setup.

A rank's assembled output is COO with duplicates, the form `MatSetValuesCOO` takes:
    [own slice | values received from rank 0 | from rank 1 | ...]   with fixed (i, j) lists.
`merge()` additionally produces the owned rows as a CSR (union pattern, explicit zeros kept,
sum in (own, source-rank) order -- deterministic) for hosts that want AIJ arrays and for the
parity tests.  This module holds index logic only; values move through torch tensors that
alias femx device buffers, and sums run in the femx_gather_sum kernel.
"""
from __future__ import annotations

import dataclasses

import numpy as np

from .mesh import petsc_ownership


@dataclasses.dataclass
class ExchangePlan:
    nranks: int
    rank: int
    row_start: int
    row_end: int
    own_lo: int                 # local value slice [own_lo, own_hi) = owned rows
    own_hi: int
    send_lo: np.ndarray         # [nranks] local value slice per destination
    send_hi: np.ndarray
    send_rows: list             # per destination: global rows / cols of the slice (setup only)
    send_cols: list
    send_frows: list = None     # per destination: global rows of the internal-force entries of the slice
    f_bounds: np.ndarray = None # [nranks+1] local dofs (pdim*node + i) by owner
    recv_counts: np.ndarray = None
    recv_fcounts: np.ndarray = None
    coo_i: np.ndarray = None    # COO of what the rank receives (source-rank order)
    coo_j: np.ndarray = None
    # receiver add over the interface rows (rows of this rank that receive contributions)
    if_rows: np.ndarray = None  # global rows, ascending
    if_rowptr: np.ndarray = None
    if_colidx: np.ndarray = None
    m_own: np.ndarray = None    # per merged entry: local value slot of this rank's own contribution, -1 = none
    m_ptr: np.ndarray = None
    m_idx: np.ndarray = None    # positions in the receive arena, (source rank, sender position) order
    f_row: np.ndarray = None    # owned global rows that receive internal-force entries
    f_ptr: np.ndarray = None
    f_idx: np.ndarray = None


def _ranges(starts: np.ndarray, lens: np.ndarray) -> np.ndarray:
    """concatenate(arange(s, s + l) for s, l in zip(starts, lens)) without a Python loop"""
    lens = np.asarray(lens, np.int64)
    tot = int(lens.sum())
    if tot == 0:
        return np.zeros(0, np.int64)
    off = np.concatenate([[0], np.cumsum(lens)[:-1]])
    return np.repeat(np.asarray(starts, np.int64) - off, lens) + np.arange(tot, dtype=np.int64)


def plan_sends(rowptr: np.ndarray, colidx: np.ndarray, nrows: int, nranks: int, rank: int, nl2g=None, pdim: int = 0,
               row_bounds=None) -> ExchangePlan:
    """Split the rank-local CSR (global row numbering, rows of absent nodes empty) by row owner.  Owners: the PETSC_DECIDE
    split (src/main.F90:154), or row_bounds[nranks + 1] when the host chose the local sizes itself (a numbering that
    follows the partition, mesh.renumber_by_partition)."""
    if row_bounds is None:
        bounds = np.array([petsc_ownership(nrows, nranks, r)[0] for r in range(nranks)] + [nrows], dtype=np.int64)
    else:
        bounds = np.asarray(row_bounds, np.int64)
        assert len(bounds) == nranks + 1 and bounds[0] == 0 and bounds[-1] == nrows and np.all(np.diff(bounds) >= 0), "row_bounds"
    lo = rowptr[bounds[:-1]].astype(np.int64)
    hi = rowptr[bounds[1:]].astype(np.int64)
    send_rows, send_cols = [], []
    for q in range(nranks):
        if q == rank or hi[q] == lo[q]:
            send_rows.append(np.zeros(0, np.int64)); send_cols.append(np.zeros(0, np.int64))
        else:
            r0, r1 = int(bounds[q]), int(bounds[q + 1])
            send_rows.append(np.repeat(np.arange(r0, r1, dtype=np.int64), np.diff(rowptr[r0:r1 + 1])))
            send_cols.append(colidx[lo[q]:hi[q]].astype(np.int64))
    ep = ExchangePlan(nranks=nranks, rank=rank, row_start=int(bounds[rank]), row_end=int(bounds[rank + 1]),
                      own_lo=int(lo[rank]), own_hi=int(hi[rank]), send_lo=lo, send_hi=hi,
                      send_rows=send_rows, send_cols=send_cols)
    if nl2g is not None:      # internal-force entries travel by local dof, pdim * node + i: the sender names the global rows
        gdof = (pdim * (np.asarray(nl2g, np.int64) - 1)[:, None] + np.arange(pdim)[None, :]).reshape(-1)
        assert np.all(np.diff(gdof) > 0), "local nodes must ascend in global id"
        ep.f_bounds = np.searchsorted(gdof, bounds).astype(np.int64)
        ep.send_frows = [gdof[ep.f_bounds[q]:ep.f_bounds[q + 1]] if q != rank else np.zeros(0, np.int64) for q in range(nranks)]
    return ep


def finish_plan(ep: ExchangePlan, rowptr: np.ndarray, colidx: np.ndarray, recv_rows: list, recv_cols: list,
                recv_frows: list | None = None) -> ExchangePlan:
    """After the one-time (row, col) exchange: COO lists of the received part and the receiver-add lists over the interface
    rows -- O(interface) setup and O(interface) work per assembly (csrc/peer.cu: peer_merge_kernel)."""
    r0, r1 = ep.row_start, ep.row_end
    ep.recv_counts = np.array([len(x) for x in recv_rows], dtype=np.int64)
    for x in recv_rows:
        assert np.all((np.asarray(x) >= r0) & (np.asarray(x) < r1)), "received a row this rank does not own"
    ep.coo_i = np.concatenate([np.asarray(x, np.int64) for x in recv_rows]) if recv_rows else np.zeros(0, np.int64)
    ep.coo_j = np.concatenate([np.asarray(x, np.int64) for x in recv_cols]) if recv_cols else np.zeros(0, np.int64)
    nrecv = len(ep.coo_i)
    ep.if_rows = np.unique(ep.coo_i)
    # this rank's own entries of those rows
    st = rowptr[ep.if_rows].astype(np.int64)
    ln = (rowptr[ep.if_rows + 1] - rowptr[ep.if_rows]).astype(np.int64)
    own_slots = _ranges(st, ln)
    own_rows = np.repeat(ep.if_rows, ln)
    own_cols = colidx[own_slots].astype(np.int64)
    ncols = int(max(ep.coo_j.max() if nrecv else 0, own_cols.max() if len(own_cols) else 0)) + 1
    key = np.concatenate([own_rows * ncols + own_cols, ep.coo_i * ncols + ep.coo_j])
    src = np.concatenate([-1 - own_slots, np.arange(nrecv, dtype=np.int64)])      # own: -1 - slot; received: arena position
    order = np.argsort(key, kind="stable")      # inside a nonzero: own first, then (source rank, sender position)
    ks, ss = key[order], src[order]
    first = np.ones(len(ks), bool)
    first[1:] = ks[1:] != ks[:-1]
    starts = np.nonzero(first)[0]
    nmer = len(starts)
    has_own = ss[starts] < 0 if nmer else np.zeros(0, bool)
    ep.m_own = np.where(has_own, -1 - ss[starts], -1).astype(np.int64) if nmer else np.zeros(0, np.int64)
    counts = np.diff(np.concatenate([starts, [len(ks)]])) - has_own.astype(np.int64) if nmer else np.zeros(0, np.int64)
    ep.m_ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    ep.m_idx = ss[ss >= 0].astype(np.int64)      # sorted order keeps groups contiguous; own entries drop out
    urow = ks[starts] // ncols if nmer else np.zeros(0, np.int64)
    ep.if_colidx = (ks[starts] % ncols).astype(np.int32) if nmer else np.zeros(0, np.int32)
    ep.if_rowptr = np.concatenate([[0], np.cumsum(np.bincount(np.searchsorted(ep.if_rows, urow), minlength=len(ep.if_rows)))]).astype(np.int64)
    # internal-force entries
    if recv_frows is not None:
        ep.recv_fcounts = np.array([len(x) for x in recv_frows], dtype=np.int64)
        fr = np.concatenate([np.asarray(x, np.int64) for x in recv_frows]) if recv_frows else np.zeros(0, np.int64)
        assert np.all((fr >= r0) & (fr < r1)), "received a force entry this rank does not own"
        fo = np.argsort(fr, kind="stable")
        frs = fr[fo]
        ff = np.ones(len(frs), bool)
        ff[1:] = frs[1:] != frs[:-1]
        fs = np.nonzero(ff)[0]
        ep.f_row = frs[fs].astype(np.int64)
        ep.f_ptr = np.concatenate([fs, [len(frs)]]).astype(np.int64)
        ep.f_idx = fo.astype(np.int64)
    return ep


def owned_csr(ep: ExchangePlan, rowptr: np.ndarray, colidx: np.ndarray, vals_local: np.ndarray, merged: np.ndarray):
    """Host view of the rank's owned rows after an assembly: the local CSR slice, with the interface rows replaced by their
    merged form (union pattern).  Used by the parity tests; a PETSc host passes the two pieces to MatSetValuesCOO instead
    (own slice with the interface rows' i set to -1, which COO ignores, plus the merged rows)."""
    r0, r1 = ep.row_start, ep.row_end
    is_if = np.zeros(r1 - r0, bool)
    is_if[ep.if_rows - r0] = True
    own_len = np.diff(rowptr[r0:r1 + 1]).astype(np.int64)
    if_len = np.zeros(r1 - r0, np.int64)
    if_len[ep.if_rows - r0] = np.diff(ep.if_rowptr)
    lens = np.where(is_if, if_len, own_len)
    rp = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    ci = np.empty(rp[-1], np.int32)
    va = np.empty(rp[-1])
    keep = ~is_if
    dst = _ranges(rp[:-1][keep], lens[keep])
    srcs = _ranges(rowptr[r0:r1][keep], lens[keep])
    ci[dst] = colidx[srcs]; va[dst] = vals_local[srcs]
    dst = _ranges(rp[:-1][is_if], lens[is_if])
    ci[dst] = ep.if_colidx; va[dst] = merged
    return rp, ci, va


def exchange_lists(ep: ExchangePlan, dist, device="cpu"):
    """One-time exchange of the (row, col) lists and of the force rows with torch.distributed (gloo or NCCL).
    Returns (recv_rows, recv_cols, recv_frows), each a list by source rank."""
    import torch
    P = ep.nranks

    def a2a(lists):
        counts = torch.tensor([len(x) for x in lists], dtype=torch.int64, device=device)
        rcounts = torch.empty(P, dtype=torch.int64, device=device)
        dist.all_to_all_single(rcounts, counts)
        rc, sc = rcounts.cpu().tolist(), counts.cpu().tolist()
        send = torch.from_numpy(np.concatenate(lists).astype(np.int64)).to(device)
        recv = torch.empty(int(sum(rc)), dtype=torch.int64, device=device)
        dist.all_to_all_single(recv, send, output_split_sizes=rc, input_split_sizes=sc)
        return np.split(recv.cpu().numpy(), np.cumsum(rc)[:-1])

    return a2a(ep.send_rows), a2a(ep.send_cols), a2a(ep.send_frows)


def peer_layout(ep: ExchangePlan, nnz: int):
    """Index side of the peer-memory exchange (csrc/peer.cu): where each source rank's slice lands in this rank's
    receive arena (source-rank order), and which local ranges go to which owner."""
    recv_voff = np.concatenate([[0], np.cumsum(ep.recv_counts)]).astype(np.int64)
    recv_foff = np.concatenate([[0], np.cumsum(ep.recv_fcounts)]).astype(np.int64)
    val_bounds = np.concatenate([ep.send_lo, [nnz]]).astype(np.int64)
    assert np.array_equal(ep.send_hi, val_bounds[1:]), "local rows must ascend with their owner"
    return dict(recv_voff=recv_voff, recv_foff=recv_foff, val_bounds=val_bounds, f_bounds=ep.f_bounds,
                recv_from=(ep.recv_counts + ep.recv_fcounts).astype(np.int64))


def exchange_offsets(lay: dict, dist, nranks: int, device="cpu"):
    """Tell every source rank where its slices start in my arena; learn where mine start in theirs.
    Returns (val_dst[P], f_dst[P], caps[P, 2]) with caps = every rank's arena capacities (values, force entries)."""
    import torch
    mine = torch.tensor(np.stack([lay["recv_voff"][:-1], lay["recv_foff"][:-1]], 1).reshape(-1), dtype=torch.int64, device=device)
    theirs = torch.empty_like(mine)
    dist.all_to_all_single(theirs, mine)
    theirs = theirs.cpu().numpy().reshape(nranks, 2)
    cap = torch.tensor([int(lay["recv_voff"][-1]), int(lay["recv_foff"][-1])], dtype=torch.int64, device=device)
    caps = [torch.empty_like(cap) for _ in range(nranks)]
    dist.all_gather(caps, cap)
    caps = np.stack([c.cpu().numpy() for c in caps])
    return theirs[:, 0].copy(), theirs[:, 1].copy(), caps


class _DevArray:
    """torch view of library-owned device memory (CUDA array interface)."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f8", "data": (int(ptr), False), "version": 3, "strides": None}


def _i64(a):
    return np.ascontiguousarray(a, np.int64)


def _set_merge(fx, ep: ExchangePlan):
    """hand the receiver-add lists to the library (interface rows only)"""
    args = [_i64(ep.m_own), _i64(ep.m_ptr), _i64(ep.m_idx), _i64(ep.f_row), _i64(ep.f_ptr), _i64(ep.f_idx)]
    p = [a.ctypes.data for a in args]
    fx._ck(fx.L.femx_peer_set_merge(fx.h, len(ep.m_own), p[0], p[1], p[2], len(ep.f_row), p[3], p[4], p[5]))


def _set_scatter(fx, ep: ExchangePlan, recv_frows: list, u_dst, recv_u):
    """reverse exchange (GetVec_U): the owned rows whose displacement each peer holds as a ghost are the rows it sends force
    entries for; they land at that peer's local dofs f_bounds_peer[my rank] + i"""
    ptr = np.concatenate([[0], np.cumsum([len(x) for x in recv_frows])]).astype(np.int64)
    rows = _i64(np.concatenate([np.asarray(x, np.int64) for x in recv_frows])) if len(recv_frows) else np.zeros(0, np.int64)
    a = [_i64(ptr), rows, _i64(u_dst), _i64(recv_u)]
    fx._ck(fx.L.femx_peer_set_scatter(fx.h, *[x.ctypes.data for x in a]))


class DistributedAssembly:
    """Per-rank driver: femx context on the local mesh + the exchange plan.

    peer = True (default on GPUs of one node): interface rows travel through femx's peer-memory arenas, the step barrier
    among neighbours and the receiver add run inside femx_assemble_* (csrc/peer.cu).  peer = False: the value slices go
    through torch.distributed all_to_all (NCCL; gloo in the CPU tests) after the call and merged_host() adds on the host.
    `vals_local` lives in a torch tensor; femx writes into it through its raw device pointer.
    deferred = True (with peer): ranks that SHARE a GPU (more ranks than GPUs) -- assemblies and scatter_u push and signal,
    complete() synchronises, meets the other ranks and runs the receiver add (femx_peer_set_deferred).  comm_device = where
    the setup collectives' tensors live ("cpu" for a gloo group; default: device, for NCCL).
    """

    def __init__(self, fx, dist, device, peer=False, row_bounds=None, deferred=False, comm_device=None):
        import torch
        self.fx, self.dist, self.device, self.torch = fx, dist, device, torch
        self.comm = comm = device if comm_device is None else comm_device
        self.deferred = False
        self.rowptr, self.colidx = fx.pattern()
        nranks, rank = dist.get_world_size(), dist.get_rank()
        m = fx.mesh
        ep = plan_sends(self.rowptr, self.colidx, fx.nrows, nranks, rank, m.nl2g, m.pdim, row_bounds)
        rr, rcs, rf = exchange_lists(ep, dist, comm)
        self._recv_frows = rf
        self.ep = finish_plan(ep, self.rowptr, self.colidx, rr, rcs, rf)
        fx._ck(fx.L.femx_set_ownership(fx.h, nranks, rank, ep.row_start, ep.row_end))
        self.vals_local = torch.empty(fx.nnz, dtype=torch.float64, device=device)
        self.recv = torch.empty(int(self.ep.recv_counts.sum()), dtype=torch.float64, device=device)
        self.send_splits = [int(h - l) if q != rank else 0 for q, (l, h) in enumerate(zip(ep.send_lo, ep.send_hi))]
        self.recv_splits = [int(x) for x in self.ep.recv_counts]
        self.n_own = self.ep.own_hi - self.ep.own_lo
        self.peer = False
        if peer:                 # all ranks or none: a rank whose IPC mapping fails takes everybody back to NCCL
            try:
                self._connect_peers()
                good = 1
            except Exception as e:   # noqa: BLE001
                import sys
                print(f"femx: peer-memory exchange unavailable on rank {rank} ({e}); using NCCL", file=sys.stderr, flush=True)
                good = 0
            t = torch.tensor([good], device=comm)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            if not int(t.item()) and good:
                self.fx._ck(self.fx.L.femx_peer_disconnect(self.fx.h))
            self.peer = bool(int(t.item()))
            if self.peer and deferred:
                fx._ck(fx.L.femx_peer_set_deferred(fx.h, 1))
                self.deferred = True

    def _connect_peers(self):
        import ctypes as C
        torch, fx, dist, ep = self.torch, self.fx, self.dist, self.ep
        P = ep.nranks
        self.lay = lay = peer_layout(ep, fx.nnz)
        val_dst, f_dst, caps = exchange_offsets(lay, dist, P, self.comm)
        handle = (C.c_char * 64)()
        fx._ck(fx.L.femx_peer_arena_create(fx.h, int(caps[ep.rank, 0]), int(caps[ep.rank, 1]), C.addressof(handle)))
        mine = torch.frombuffer(bytearray(handle.raw), dtype=torch.uint8).to(self.comm)
        alls = [torch.empty_like(mine) for _ in range(P)]
        dist.all_gather(alls, mine)
        handles = b"".join(bytes(t.cpu().numpy().tobytes()) for t in alls)
        args = [_i64(caps[:, 0]), _i64(caps[:, 1]), _i64(lay["val_bounds"]), _i64(val_dst), _i64(lay["f_bounds"]), _i64(f_dst), _i64(lay["recv_from"])]
        fx._ck(fx.L.femx_peer_connect(fx.h, P, ep.rank, handles, *[a.ctypes.data for a in args]))
        _set_merge(fx, ep)
        self.n_recv_f = int(caps[ep.rank, 1])
        # GetVec_U: every rank tells its peers where (in its local dof numbering) the block they own starts, and its pdim * nlnds
        mine = torch.tensor(np.stack([lay["f_bounds"][:-1], np.full(P, fx.mesh.pdim * fx.mesh.nlnds)], 1).reshape(-1), dtype=torch.int64, device=self.comm)
        theirs = torch.empty_like(mine)
        dist.all_to_all_single(theirs, mine)
        theirs = theirs.cpu().numpy().reshape(P, 2)
        _set_scatter(fx, ep, self._recv_frows, theirs[:, 0], theirs[:, 1])

    def scatter_u(self, u_owned, u_local):
        """GetVec_U (src/m_global.F90:549-564): u_owned = this rank's slice of the global vector (device tensor),
        u_local = the ghosted local vector [pdim * nlnds] (device tensor, filled)."""
        import ctypes as C
        self.fx._ck(self.fx.L.femx_scatter_u(self.fx.h, C.c_void_p(u_owned.data_ptr()), C.c_void_p(u_local.data_ptr()), 1))
        return u_local

    def complete(self):
        """deferred peers: finish the outstanding assembly / scatter_u once every rank has issued its own"""
        self.torch.cuda.synchronize(self.device)
        self.dist.barrier()
        self.fx._ck(self.fx.L.femx_peer_complete(self.fx.h))
        self.dist.barrier()      # nobody starts the next step's pushes into a buffer a peer is still adding from

    def peer_received(self):
        """(values, force entries) the last assembly received, as torch views of the arena."""
        import ctypes as C
        pv, pf = C.c_void_p(), C.c_void_p()
        self.fx._ck(self.fx.L.femx_peer_recv(self.fx.h, C.byref(pv), C.byref(pf)))
        n = int(self.ep.recv_counts.sum())
        as_t = lambda p, k: (self.torch.as_tensor(_DevArray(p.value, k), device=self.device) if k > 0
                             else self.torch.empty(0, dtype=self.torch.float64, device=self.device))
        return as_t(pv, n), as_t(pf, self.n_recv_f)

    def merged(self):
        """Merged interface-row values of the last peer-mode assembly (device tensor view, union pattern ep.if_*)."""
        import ctypes as C
        p, n = C.c_void_p(), C.c_int64()
        self.fx._ck(self.fx.L.femx_peer_merged(self.fx.h, C.byref(p), C.byref(n)))
        if n.value == 0:
            return self.torch.empty(0, dtype=self.torch.float64, device=self.device)
        return self.torch.as_tensor(_DevArray(p.value, n.value), device=self.device)

    def close(self):
        if self.peer:
            self.fx._ck(self.fx.L.femx_peer_disconnect(self.fx.h))
            self.dist.barrier()  # nobody frees an arena a peer still maps
            self.peer = False

    def exchange(self):
        """NCCL / gloo path: ship the interface-row values (contiguous slices of vals_local, no pack kernel); returns the
        received values ordered by source rank.  In peer mode the kernel has already delivered and merged them."""
        ep = self.ep
        if self.peer:
            self.recv = self.peer_received()[0]
            return self.recv
        ins = [self.vals_local[int(ep.send_lo[q]):int(ep.send_hi[q])] if q != ep.rank else self.vals_local[:0]
               for q in range(ep.nranks)]
        offs = np.concatenate([[0], np.cumsum(self.recv_splits)])
        outs = [self.recv[int(offs[q]):int(offs[q + 1])] for q in range(ep.nranks)]
        self.dist.all_to_all(outs, ins)
        return self.recv

    def merged_host(self, vals_local: np.ndarray, recv: np.ndarray) -> np.ndarray:
        """The receiver add on the host, same lists and order as peer_merge_kernel (NCCL / gloo path, CPU tests)."""
        ep = self.ep
        out = np.where(ep.m_own >= 0, vals_local[np.maximum(ep.m_own, 0)], 0.0)
        cnt = np.diff(ep.m_ptr)
        for k in range(int(cnt.max()) if len(cnt) else 0):          # k-th received contribution of every entry, in list order
            sel = np.nonzero(cnt > k)[0]
            out[sel] = out[sel] + recv[ep.m_idx[ep.m_ptr[sel] + k]]
        return out

    def owned_csr(self, vals_local: np.ndarray, merged: np.ndarray):
        return owned_csr(self.ep, self.rowptr, self.colidx, vals_local, merged)


class LocalRanks:
    """N ranks as N femx contexts of THIS process on ONE device (femx_peer_connect_local): the whole multi-rank data
    path -- per-owner value ranges, arena layout, remote stores, step flags, receiver add -- on a single GPU.  The ranks'
    kernels run one after another, so assemblies push and signal only and femx_peer_complete finishes every rank once all
    have assembled (a kernel that waited for a peer not yet launched would never finish)."""

    def __init__(self, fxs: list, row_bounds=None):
        import ctypes as C
        self.fxs, P = fxs, len(fxs)
        self.P = P
        pats = [fx.pattern() for fx in fxs]
        eps = [plan_sends(pats[r][0], pats[r][1], fxs[r].nrows, P, r, fxs[r].mesh.nl2g, fxs[r].mesh.pdim, row_bounds) for r in range(P)]
        self.eps = [finish_plan(eps[r], pats[r][0], pats[r][1], [eps[s].send_rows[r] for s in range(P)], [eps[s].send_cols[r] for s in range(P)],
                                [eps[s].send_frows[r] for s in range(P)]) for r in range(P)]
        self.pats = pats
        lays = [peer_layout(self.eps[r], fxs[r].nnz) for r in range(P)]
        caps = np.array([[int(l["recv_voff"][-1]), int(l["recv_foff"][-1])] for l in lays], np.int64)
        for r, fx in enumerate(fxs):
            handle = (C.c_char * 64)()
            fx._ck(fx.L.femx_peer_arena_create(fx.h, int(caps[r, 0]), int(caps[r, 1]), C.addressof(handle)))
        ptrs = (C.c_void_p * P)(*[fx.h for fx in fxs])
        for r, fx in enumerate(fxs):
            val_dst = _i64([lays[q]["recv_voff"][r] for q in range(P)])      # where my slice starts in rank q's arena
            f_dst = _i64([lays[q]["recv_foff"][r] for q in range(P)])
            args = [_i64(caps[:, 0]), _i64(caps[:, 1]), _i64(lays[r]["val_bounds"]), val_dst, _i64(lays[r]["f_bounds"]), f_dst, _i64(lays[r]["recv_from"])]
            fx._ck(fx.L.femx_set_ownership(fx.h, P, r, self.eps[r].row_start, self.eps[r].row_end))
            fx._ck(fx.L.femx_peer_connect_local(fx.h, P, r, C.addressof(ptrs), *[a.ctypes.data for a in args]))
            _set_merge(fx, self.eps[r])
            _set_scatter(fx, self.eps[r], [eps[s].send_frows[r] for s in range(P)], [lays[q]["f_bounds"][r] for q in range(P)],
                         [fxs[q].mesh.pdim * fxs[q].mesh.nlnds for q in range(P)])

    def scatter_u(self, u_global: np.ndarray):
        """GetVec_U on every rank: each gets its slice of the global vector and returns its ghosted local vector (host)."""
        from . import api
        bufs = []
        for r, fx in enumerate(self.fxs):
            ep = self.eps[r]
            d_o = api.DeviceBuffer(fx, max(8, 8 * (ep.row_end - ep.row_start))).upload(np.ascontiguousarray(u_global[ep.row_start:ep.row_end]))
            d_l = api.DeviceBuffer(fx, 8 * fx.mesh.pdim * fx.mesh.nlnds)
            fx._ck(fx.L.femx_scatter_u(fx.h, d_o.ptr, d_l.ptr, api.DEVICE))        # push + signal
            bufs.append((d_o, d_l))
        out = []
        for r, fx in enumerate(self.fxs):
            fx._ck(fx.L.femx_peer_complete(fx.h))                                  # flags + local gather
        for r, fx in enumerate(self.fxs):
            out.append(bufs[r][1].download().reshape(-1, fx.mesh.pdim))
            for b in bufs[r]:
                b.free()
        return out

    def assemble_K_fint(self, us: list):
        """One collective assembly.  Returns per rank (rowptr, colidx, vals) of the owned rows (host) and the global internal
        force vector summed over the ranks' owned entries."""
        import ctypes as C
        from . import api
        P = self.P
        bufs = []
        for r, fx in enumerate(self.fxs):
            d_u = api.DeviceBuffer(fx, us[r].nbytes).upload(np.ascontiguousarray(us[r]))
            d_v = api.DeviceBuffer(fx, 8 * fx.nnz)
            d_f = api.DeviceBuffer(fx, 8 * fx.nrows)
            fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, d_v.ptr, d_f.ptr, api.DEVICE))      # kernel + push + signal
            bufs.append((d_u, d_v, d_f))
        out, f_tot = [], np.zeros(self.fxs[0].nrows)
        for r, fx in enumerate(self.fxs):
            fx._ck(fx.L.femx_peer_complete(fx.h))                                              # flags + receiver add
        for r, fx in enumerate(self.fxs):
            ep = self.eps[r]
            d_u, d_v, d_f = bufs[r]
            p, n = C.c_void_p(), C.c_int64()
            fx._ck(fx.L.femx_peer_merged(fx.h, C.byref(p), C.byref(n)))
            merged = np.empty(n.value)
            if n.value:
                fx._ck(fx.L.femx_memcpy(fx.h, C.c_void_p(merged.ctypes.data), p, merged.nbytes, api.HOST, api.DEVICE))
            out.append(owned_csr(ep, self.pats[r][0], self.pats[r][1], d_v.download(), merged))
            f = d_f.download()
            f_tot[ep.row_start:ep.row_end] = f[ep.row_start:ep.row_end]
            for b in bufs[r]:
                b.free()
        return out, f_tot

    def close(self):
        for fx in self.fxs:
            fx._ck(fx.L.femx_peer_disconnect(fx.h))


# --------------------------------------------------------------------------- nodal stress across ranks
def shared_node_lists(nl2g: np.ndarray, all_nl2g: list, rank: int):
    """For every peer q: positions (in MY local numbering) of the nodes rank q holds too, in ascending global id --
    the same order on both sides, so slot i of what I send is slot i of what q expects from me."""
    mine = np.asarray(nl2g, np.int64)
    out = []
    for q, other in enumerate(all_nl2g):
        if q == rank:
            out.append(np.zeros(0, np.int64))
        else:
            common = np.intersect1d(mine, np.asarray(other, np.int64), assume_unique=True)
            out.append(np.searchsorted(mine, common))
    return out


def combine_partials(own: np.ndarray, recv: list, shared: list, rank: int):
    """Sum of the ranks' partials in ascending RANK order (deterministic, identical on every rank that shares the
    node): contributions of lower ranks, then mine, then higher ranks.  Works on numpy arrays or torch tensors."""
    total = own * 0
    for q in range(len(shared)):
        if q == rank:
            total = total + own
        elif len(shared[q]):
            total[shared[q]] = total[shared[q]] + recv[q]
    return total


class DistributedStress:
    """RecoverStress / RecoverNodalStress + the cross-rank nodal average (src/main.F90:253-291): every rank computes
    the sums of its own elements (femx_stress_partial); nodes shared with other ranks get the peers' partial sums
    (torch.distributed all_to_all over NCCL; gloo in the CPU tests) added in rank order and are divided by the global
    count -- what the reference obtains from the dense Mat_Stress assembly and MPI_AllReduce of count_node every step.
    The counts are mesh constants and are exchanged once."""

    def __init__(self, fx, dist, device):
        import torch
        self.fx, self.dist, self.device, self.torch = fx, dist, device, torch
        P, rank = dist.get_world_size(), dist.get_rank()
        self.P, self.rank = P, rank
        m = fx.mesh
        self.cp = m.cpdim
        n_all = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(P)]
        dist.all_gather(n_all, torch.tensor([m.nlnds], dtype=torch.int64, device=device))
        sizes = [int(t.item()) for t in n_all]
        pad = max(sizes)
        buf = torch.zeros(pad, dtype=torch.int64, device=device)
        buf[:m.nlnds] = torch.from_numpy(np.asarray(m.nl2g, np.int64)).to(device)
        alls = [torch.empty_like(buf) for _ in range(P)]
        dist.all_gather(alls, buf)
        all_nl2g = [alls[q][:sizes[q]].cpu().numpy() for q in range(P)]
        self.shared = shared_node_lists(m.nl2g, all_nl2g, rank)
        self.shared_dev = [torch.from_numpy(s).to(device) for s in self.shared]
        self.sums = torch.empty(m.nlnds * self.cp, dtype=torch.float64, device=device)
        counts = torch.empty(m.nlnds, dtype=torch.int32, device=device)
        self._counts_local = counts
        self.count_global = None

    def _exchange(self, rows):
        """rows [nlnds, k] tensor of partials -> list of what each peer holds for the nodes we share."""
        torch = self.torch
        ins = [rows[self.shared_dev[q]].contiguous() for q in range(self.P)]
        outs = [torch.empty_like(t) for t in ins]
        self.dist.all_to_all(outs, ins)
        return outs

    def stress(self, d_u_ptr):
        """Averaged nodal stress [nlnds, cp] of this rank's local nodes (device tensor); d_u_ptr = device pointer of uu_local."""
        import ctypes as C
        torch, fx = self.torch, self.fx
        first = self.count_global is None
        fx._ck(fx.L.femx_stress_partial(fx.h, d_u_ptr, C.c_void_p(self.sums.data_ptr()),
                                        C.c_void_p(self._counts_local.data_ptr()) if first else None, 1))
        if first:
            cl = self._counts_local.to(torch.float64).reshape(-1, 1)
            self.count_global = combine_partials(cl, self._exchange(cl), self.shared_dev, self.rank)
        part = self.sums.reshape(-1, self.cp)
        total = combine_partials(part, self._exchange(part), self.shared_dev, self.rank)
        return total / self.count_global
