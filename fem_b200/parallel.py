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
    recv_counts: np.ndarray = None
    coo_i: np.ndarray = None    # COO of the rank's output (own slice first, then per source rank)
    coo_j: np.ndarray = None
    # merged owned CSR
    rowptr: np.ndarray = None   # [row_end-row_start+1]
    colidx: np.ndarray = None
    merge_ptr: np.ndarray = None
    merge_idx: np.ndarray = None


def plan_sends(rowptr: np.ndarray, colidx: np.ndarray, nrows: int, nranks: int, rank: int) -> ExchangePlan:
    """Split the rank-local CSR (global row numbering, rows of absent nodes empty) by row owner."""
    bounds = np.array([petsc_ownership(nrows, nranks, r)[0] for r in range(nranks)] + [nrows], dtype=np.int64)
    lo = rowptr[bounds[:-1]].astype(np.int64)
    hi = rowptr[bounds[1:]].astype(np.int64)
    rows_of = lambda r0, r1: np.repeat(np.arange(r0, r1, dtype=np.int64), np.diff(rowptr[r0:r1 + 1]))
    send_rows, send_cols = [], []
    for q in range(nranks):
        if q == rank:
            send_rows.append(np.zeros(0, np.int64)); send_cols.append(np.zeros(0, np.int64))
        else:
            send_rows.append(rows_of(bounds[q], bounds[q + 1]))
            send_cols.append(colidx[lo[q]:hi[q]].astype(np.int64))
    return ExchangePlan(nranks=nranks, rank=rank, row_start=int(bounds[rank]), row_end=int(bounds[rank + 1]),
                        own_lo=int(lo[rank]), own_hi=int(hi[rank]), send_lo=lo, send_hi=hi,
                        send_rows=send_rows, send_cols=send_cols)


def finish_plan(ep: ExchangePlan, rowptr: np.ndarray, colidx: np.ndarray, recv_rows: list, recv_cols: list,
                with_merge: bool = True) -> ExchangePlan:
    """After the one-time (row, col) exchange: COO lists and the merged owned CSR."""
    r0, r1 = ep.row_start, ep.row_end
    ep.recv_counts = np.array([len(x) for x in recv_rows], dtype=np.int64)
    for x in recv_rows:
        assert np.all((np.asarray(x) >= r0) & (np.asarray(x) < r1)), "received a row this rank does not own"
    if not with_merge:          # COO lists of the received part only (the own part is the local CSR slice)
        ep.coo_i = np.concatenate([np.asarray(x, np.int64) for x in recv_rows]) if recv_rows else np.zeros(0, np.int64)
        ep.coo_j = np.concatenate([np.asarray(x, np.int64) for x in recv_cols]) if recv_cols else np.zeros(0, np.int64)
        return ep
    own_rows = np.repeat(np.arange(r0, r1, dtype=np.int64), np.diff(rowptr[r0:r1 + 1]))
    own_cols = colidx[ep.own_lo:ep.own_hi].astype(np.int64)
    ep.coo_i = np.concatenate([own_rows] + [np.asarray(x, np.int64) for x in recv_rows])
    ep.coo_j = np.concatenate([own_cols] + [np.asarray(x, np.int64) for x in recv_cols])
    # union pattern; stable sort keeps (own, source rank, sender position) order inside each nonzero
    ncols = int(ep.coo_j.max()) + 1 if len(ep.coo_j) else 1
    key = (ep.coo_i - r0) * ncols + ep.coo_j
    order = np.argsort(key, kind="stable")
    ks = key[order]
    first = np.ones(len(ks), bool)
    first[1:] = ks[1:] != ks[:-1]
    starts = np.nonzero(first)[0]
    ep.merge_ptr = np.concatenate([starts, [len(ks)]]).astype(np.int64)
    ep.merge_idx = order.astype(np.int64)
    urow = (ks[starts] // ncols).astype(np.int64)
    ep.colidx = (ks[starts] % ncols).astype(np.int32)
    ep.rowptr = np.zeros(r1 - r0 + 1, np.int64)
    np.add.at(ep.rowptr, urow + 1, 1)
    np.cumsum(ep.rowptr, out=ep.rowptr)
    return ep


def exchange_lists(ep: ExchangePlan, dist, device="cpu"):
    """One-time exchange of the (row, col) lists with torch.distributed (gloo or NCCL)."""
    import torch
    P = ep.nranks
    counts = torch.tensor([len(x) for x in ep.send_rows], dtype=torch.int64, device=device)
    rcounts = torch.empty(P, dtype=torch.int64, device=device)
    dist.all_to_all_single(rcounts, counts)
    rc = rcounts.cpu().tolist()
    sc = counts.cpu().tolist()
    out = []
    for lists in (ep.send_rows, ep.send_cols):
        send = torch.from_numpy(np.concatenate(lists)).to(device)
        recv = torch.empty(int(sum(rc)), dtype=torch.int64, device=device)
        dist.all_to_all_single(recv, send, output_split_sizes=rc, input_split_sizes=sc)
        out.append(np.split(recv.cpu().numpy(), np.cumsum(rc)[:-1]))
    return out[0], out[1]


def peer_layout(ep: ExchangePlan, recv_rows: list, nl2g: np.ndarray, pdim: int, nrows: int, nnz: int):
    """Index side of the peer-memory exchange (csrc/peer.cu): where each source rank's slice lands in this
    rank's receive arena, and which local ranges go to which owner.  Returns a dict of int64 arrays:
      recv_voff / recv_foff [P+1]  start of source s's values / force entries in the arena (source-rank order),
      f_rows[s]                    global rows of the force entries source s sends (its local rows in my range,
                                   ascending = the unique rows of its matrix list, every row has a diagonal),
      val_bounds / f_bounds [P+1]  local value offsets / local dofs (pdim*node + i) by owner."""
    P = ep.nranks
    f_rows = [np.unique(np.asarray(r, np.int64)) for r in recv_rows]
    recv_voff = np.concatenate([[0], np.cumsum([len(r) for r in recv_rows])]).astype(np.int64)
    recv_foff = np.concatenate([[0], np.cumsum([len(r) for r in f_rows])]).astype(np.int64)
    val_bounds = np.concatenate([ep.send_lo, [nnz]]).astype(np.int64)
    assert np.array_equal(ep.send_hi, val_bounds[1:]), "local rows must ascend with their owner"
    bounds = np.array([petsc_ownership(nrows, P, r)[0] for r in range(P)] + [nrows], dtype=np.int64)
    gdof = (pdim * (np.asarray(nl2g, np.int64) - 1)[:, None] + np.arange(pdim)[None, :]).reshape(-1)
    assert np.all(np.diff(gdof) > 0), "local nodes must ascend in global id"
    f_bounds = np.searchsorted(gdof, bounds).astype(np.int64)
    return dict(recv_voff=recv_voff, recv_foff=recv_foff, f_rows=f_rows, val_bounds=val_bounds, f_bounds=f_bounds)


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


class DistributedAssembly:
    """Per-rank driver: femx context on the local mesh + the exchange plan.

    `vals_local` lives in a torch tensor so that torch.distributed (NCCL) can send slices of it;
    femx writes into it through its raw device pointer.
    """

    def __init__(self, fx, dist, device, with_merge=True, peer=False):
        import torch
        self.fx, self.dist, self.device, self.torch = fx, dist, device, torch
        rowptr, colidx = fx.pattern()
        nranks, rank = dist.get_world_size(), dist.get_rank()
        ep = plan_sends(rowptr, colidx, fx.nrows, nranks, rank)
        rr, rcs = exchange_lists(ep, dist, device)
        self.ep = finish_plan(ep, rowptr, colidx, rr, rcs, with_merge)
        self.vals_local = torch.empty(fx.nnz, dtype=torch.float64, device=device)
        self.recv = torch.empty(int(self.ep.recv_counts.sum()), dtype=torch.float64, device=device)
        self.send_splits = [int(h - l) if q != rank else 0 for q, (l, h) in enumerate(zip(ep.send_lo, ep.send_hi))]
        self.recv_splits = [int(x) for x in self.ep.recv_counts]
        self.n_own = self.ep.own_hi - self.ep.own_lo
        self._merge_dev = None
        self.peer = False
        if peer:                 # all ranks or none: a rank whose IPC mapping fails takes everybody back to NCCL
            try:
                self._connect_peers(rr)
                good = 1
            except Exception as e:   # noqa: BLE001
                import sys
                print(f"femx: peer-memory exchange unavailable on rank {rank} ({e}); using NCCL", file=sys.stderr, flush=True)
                good = 0
            t = torch.tensor([good], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            if not int(t.item()) and good:
                self.fx._ck(self.fx.L.femx_peer_disconnect(self.fx.h))
            self.peer = bool(int(t.item()))

    def _connect_peers(self, recv_rows):
        """Peer-memory mode: the assembly kernel stores interface rows into the owner's arena (csrc/peer.cu);
        exchange() then only looks up where the last assembly's receive buffer is."""
        import ctypes as C
        torch, fx, dist, ep = self.torch, self.fx, self.dist, self.ep
        P = ep.nranks
        m = fx.mesh
        self.lay = lay = peer_layout(ep, recv_rows, m.nl2g, m.pdim, fx.nrows, fx.nnz)
        val_dst, f_dst, caps = exchange_offsets(lay, dist, P, self.device)
        handle = (C.c_char * 64)()
        fx._ck(fx.L.femx_peer_arena_create(fx.h, int(caps[ep.rank, 0]), int(caps[ep.rank, 1]), C.addressof(handle)))
        mine = torch.frombuffer(bytearray(handle.raw), dtype=torch.uint8).to(self.device)
        alls = [torch.empty_like(mine) for _ in range(P)]
        dist.all_gather(alls, mine)
        handles = b"".join(bytes(t.cpu().numpy().tobytes()) for t in alls)
        i64 = lambda a: np.ascontiguousarray(a, np.int64)
        args = [i64(caps[:, 0]), i64(caps[:, 1]), i64(lay["val_bounds"]), i64(val_dst), i64(lay["f_bounds"]), i64(f_dst)]
        fx._ck(fx.L.femx_peer_connect(fx.h, P, ep.rank, handles, *[a.ctypes.data for a in args]))
        self._peer_keep = args
        self.n_recv_f = int(caps[ep.rank, 1])

    def peer_received(self):
        """(values, force entries) the last assembly received, as torch views of the arena."""
        import ctypes as C
        pv, pf = C.c_void_p(), C.c_void_p()
        self.fx._ck(self.fx.L.femx_peer_recv(self.fx.h, C.byref(pv), C.byref(pf)))
        n = int(self.ep.recv_counts.sum())
        as_t = lambda p, k: (self.torch.as_tensor(_DevArray(p.value, k), device=self.device) if k > 0
                             else self.torch.empty(0, dtype=self.torch.float64, device=self.device))
        return as_t(pv, n), as_t(pf, self.n_recv_f)

    def owned_fint(self, d_f):
        """Internal force of the owned rows after a peer-mode assembly: this rank's partial plus the entries
        each source rank stored into the arena, added in source-rank order (deterministic)."""
        torch, ep = self.torch, self.ep
        out = d_f[ep.row_start:ep.row_end].clone()
        fr = self.peer_received()[1]
        for s, rows in enumerate(self.lay["f_rows"]):
            if len(rows):
                o = int(self.lay["recv_foff"][s])
                idx = torch.from_numpy(rows - ep.row_start).to(self.device)
                out.index_add_(0, idx, fr[o:o + len(rows)])
        return out

    def close(self):
        if self.peer:
            self.fx._ck(self.fx.L.femx_peer_disconnect(self.fx.h))
            self.dist.barrier()  # nobody frees an arena a peer still maps
            self.peer = False

    def exchange(self):
        """Ship the interface-row values (contiguous slices of vals_local, no pack kernel);
        returns (own slice view, received values ordered by source rank)."""
        ep = self.ep
        if self.peer:            # the kernel already delivered the slices; the step barrier ran inside the assemble call
            self.recv = self.peer_received()[0]
            return self.vals_local[ep.own_lo:ep.own_hi], self.recv
        ins = [self.vals_local[int(ep.send_lo[q]):int(ep.send_hi[q])] if q != ep.rank else self.vals_local[:0]
               for q in range(ep.nranks)]
        offs = np.concatenate([[0], np.cumsum(self.recv_splits)])
        outs = [self.recv[int(offs[q]):int(offs[q + 1])] for q in range(ep.nranks)]
        self.dist.all_to_all(outs, ins)
        return self.vals_local[ep.own_lo:ep.own_hi], self.recv

    def merged(self):
        """Owned rows as CSR values (union pattern), summed in (own, source rank) order on the device."""
        import ctypes as C
        torch, ep, fx = self.torch, self.ep, self.fx
        if self._merge_dev is None:
            self._merge_dev = (torch.from_numpy(ep.merge_ptr).to(self.device), torch.from_numpy(ep.merge_idx).to(self.device),
                               torch.empty(len(ep.colidx), dtype=torch.float64, device=self.device))
        ptr, idx, out = self._merge_dev
        own = self.vals_local[ep.own_lo:ep.own_hi]
        torch.cuda.synchronize()
        fx._ck(fx.L.femx_gather_sum(fx.h, C.c_void_p(own.data_ptr()), self.n_own, C.c_void_p(self.recv.data_ptr()),
                                    C.c_void_p(ptr.data_ptr()), C.c_void_p(idx.data_ptr()), len(ep.colidx),
                                    C.c_void_p(out.data_ptr())))
        return out


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
