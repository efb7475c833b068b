#!/usr/bin/env python
"""Multi-GPU parity check (run under torchrun, one rank per GPU):
    torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_parity.py
Every rank assembles its slab with the CUDA path; interface rows travel either inside the assembly kernel
(peer-memory stores over NVLink, csrc/peer.cu) or over NCCL after it, and the merged owned rows are compared
with the CPU oracle's 1-rank global matrix (pattern bit-exact, values 1e-12), twice per mode so that both
receive buffers of the peer arena are exercised.
FEMX_SAME_DEVICE=1: the ranks are PROCESSES that share GPU 0 (more ranks than GPUs): gloo for the setup exchange, the same
IPC-mapped arenas, deferred completion (femx_peer_set_deferred + femx_peer_complete after a host barrier) because kernels
of different processes are time-sliced; the NCCL mode is left out (NCCL refuses two ranks on one device)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import api, mesh as fm, meshgen as mg, parallel as par  # noqa: E402
from oracle import pyoracle as orc  # noqa: E402


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    same = os.environ.get("FEMX_SAME_DEVICE") == "1"
    if same:
        lr = 0
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    if same:
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=dev)
    comm = "cpu" if same else dev
    ok = True
    import ctypes as C
    cases = [(name, g, peer, rep) for name, g in (("hex", mg.jitter(mg.hex_cube(12, clamp=False))), ("tet", mg.jitter(mg.tet_cube(6))),
                                                  ("quad", mg.jitter(mg.quad_rect(40, 31))))
             for peer in ((1, 2) if same else (0, 1, 2)) for rep in ((0, 1, 2) if peer else (0,))]   # 0 NCCL, 1 push kernel, 2 stores fused into the assembly kernel
    fx = da = None
    for name, g, peer, rep in cases:
        if rep == 0:
            if fx is not None:
                da.close(); fx.close()
            loc = fm.localize(g, mg.slab_partition(g, world), rank)
            fx = api.Femx(device=lr, peer_in_kernel=peer == 2)
            fx.set_mesh(loc).build_plan()
            da = par.DistributedAssembly(fx, dist, dev, peer=peer > 0, deferred=same, comm_device=comm)
            assert not same or (da.peer and da.deferred), "same-device ranks need the IPC arenas"
        u = mg.displacement_field(loc)
        d_u = torch.from_numpy(u.reshape(-1)).to(dev)
        d_f = torch.zeros(fx.nrows, dtype=torch.float64, device=dev)
        da.vals_local.fill_(float("nan"))
        torch.cuda.synchronize()
        fx._ck(fx.L.femx_assemble_K_fint(fx.h, C.c_void_p(d_u.data_ptr()), C.c_void_p(da.vals_local.data_ptr()),
                                         C.c_void_p(d_f.data_ptr()), api.DEVICE))
        if same:
            da.complete()
        recv = da.exchange()
        vl = da.vals_local.cpu().numpy()
        if peer:                                   # the receiver add ran inside the assemble call (interface rows only)
            merged = da.merged().cpu().numpy()
            d_f2 = torch.zeros_like(d_f)
            d_f2[da.ep.row_start:da.ep.row_end] = d_f[da.ep.row_start:da.ep.row_end]     # owned rows: own partial + received, added by femx
            d_f = d_f2
        else:
            merged = da.merged_host(vl, recv.cpu().numpy())
        rp_o, ci_o, va_o = da.owned_csr(vl, merged)
        if same:
            d_f = d_f.cpu()
        dist.all_reduce(d_f)                       # residual: the ranks' owned entries (peer) / partial vectors (NCCL) sum to the global vector
        og = orc.Mesh(g)
        rp, ci, va = orc.assemble_K(og).export(og.ndof)
        r0, r1 = da.ep.row_start, da.ep.row_end
        ref = va[rp[r0]:rp[r1]]
        good = (np.array_equal(rp_o, rp[r0:r1 + 1] - rp[r0]) and np.array_equal(ci_o, ci[rp[r0]:rp[r1]])
                and np.all(np.abs(va_o - ref) <= 1e-12 * np.maximum(np.abs(ref), np.abs(va).max())))
        ug = np.zeros((g.nnds, g.pdim)); ug[loc.nl2g - 1] = u
        f_ref = orc.fint(og, mg.displacement_field(g).reshape(-1))
        f = d_f.cpu().numpy()
        good = good and np.all(np.abs(f - f_ref) <= 1e-12 * np.maximum(np.abs(f_ref), np.abs(f_ref).max()))
        if peer == 1:                              # GetVec_U through the arenas: owned slice of the global vector -> ghosted local vector
            ug_all = torch.from_numpy(mg.displacement_field(g).reshape(-1) + rep).to(dev)
            ul = da.scatter_u(ug_all[da.ep.row_start:da.ep.row_end].contiguous(), torch.empty(loc.pdim * loc.nlnds, dtype=torch.float64, device=dev))
            if same:
                da.complete()
            want_u = ug_all.cpu().numpy().reshape(-1, g.pdim)[loc.nl2g - 1].reshape(-1)
            good = good and np.array_equal(ul.cpu().numpy(), want_u)
        if rep == 0 and peer == 0:                 # nodal stress across ranks (femx_stress_partial + DistributedStress)
            ds = par.DistributedStress(fx, dist, dev)
            s_loc = ds.stress(C.c_void_p(d_u.data_ptr())).cpu().numpy()
            s_loc2 = ds.stress(C.c_void_p(d_u.data_ptr())).cpu().numpy()
            s_ref, _ = orc.stress(og, mg.displacement_field(g).reshape(-1))
            want = s_ref[loc.nl2g - 1]
            sgood = np.all(np.abs(s_loc - want) <= 1e-10 * np.maximum(np.abs(want), np.abs(s_ref).max())) and np.array_equal(s_loc, s_loc2)
            print(f"[rank {rank}] {name} nodal stress over {world} ranks {'OK' if sgood else 'MISMATCH'}", flush=True)
            good = good and bool(sgood)
        print(f"[rank {rank}] {name} {('nccl', 'peer-push', 'peer-fused')[peer]} #{rep}: owned rows [{r0},{r1}) nnz {len(va_o)} interface nnz {len(merged)} recv {int(da.ep.recv_counts.sum())} "
              f"{'OK' if good else 'MISMATCH'}", flush=True)
        ok = ok and bool(good)
    da.close(); fx.close()
    t = torch.tensor([1 if ok else 0], device=comm)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("MGPU PARITY", "OK" if int(t.item()) else "FAILED", flush=True)
    dist.destroy_process_group()
    return 0 if int(t.item()) else 1


if __name__ == "__main__":
    sys.exit(main())
