#!/usr/bin/env python
"""Strong-scaling lines for the BASELINE.json configurations that bench.py (weak scaling of the 100^3 headline cube)
does not cover.  Run under torchrun, one rank per GPU:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/scale_bench.py --config c5|c3 [--steps K --warmup W --exchange peer|nccl]
  c5  hex8 200^3 (8,000,000 elements, 24.4 M dof, 1.95 G nonzeros) cut into N z-slabs, built rank-locally
  c3  tet4 94^3 x 6 (4,983,504 elements), K + f_int (+ --stress: nodal-averaged stress), --partition
        slab              N slabs of the global mesh in element order (mesh.localize)
        metis             METIS_PartMeshNodal as src/m_io.F90:137 calls it, the mesh file's node numbering, PETSC_DECIDE rows
                          (the reference's situation: most rows a rank assembles belong to other ranks)
        metis-renumbered  the same partition, node ids following the parts (mesh.renumber_by_partition), rows = the parts' ranges
c5 rows follow the PETSC_DECIDE split; interface rows travel through the peer-memory push kernel, the neighbours' step barrier and
the receiver add inside the timed call (csrc/peer.cu), or NCCL after it.  Time = device time of femx_assemble_K_fint per step
(CUDA events on the library's stream), max over ranks.  One JSON line from rank 0."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fem_b200 import api, mesh as fm, meshgen as mg, parallel as par  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", choices=("c5", "c3"), default="c5")
    ap.add_argument("--edge", type=int, default=0, help="cube edge (cells): default 200 for c5, 94 for c3")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--exchange", choices=("peer", "nccl"), default="peer")
    ap.add_argument("--partition", choices=("slab", "metis", "metis-renumbered"), default="slab",
                    help="c3: element partition -- slabs, or METIS_PartMeshNodal called the way src/m_io.F90:137 calls it (rank 0 partitions, epart is "
                         "broadcast); metis-renumbered: the same partition with a node numbering that follows it and the parts' own row ranges "
                         "(mesh.renumber_by_partition) instead of the reference's natural numbering + PETSC_DECIDE")
    ap.add_argument("--stress", action="store_true", help="also time the nodal-averaged stress recovery (wall clock per call)")
    args = ap.parse_args()
    rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    t0 = time.time()
    if args.config == "c5":
        n = args.edge or 200
        z = [n * r // world for r in range(world + 1)]
        m = mg.hex_block(n, n, z[rank + 1] - z[rank], z0=z[rank], nz_total=n)
        total, label = n ** 3, f"C5 hex8 {n}^3, {world} z-slab(s)"
    else:
        n = args.edge or 94
        g = mg.jitter(mg.tet_cube(n))
        total, label = g.nels, f"C3 tet4 {n}^3x6, {world} slab(s)"
        row_bounds = None
        if world > 1 and args.partition.startswith("metis"):
            ep_t = torch.zeros(g.nels, dtype=torch.int32, device=dev)
            if rank == 0:
                tm = time.time()
                epart, cut = fm.metis_partition(g, world)
                label = f"C3 tet4 {n}^3x6, METIS_PartMeshNodal into {world} parts (edge cut {cut}, {time.time() - tm:.1f} s on one host core)"
                ep_t.copy_(torch.from_numpy(epart))
            dist.broadcast(ep_t, 0)
            epart = ep_t.cpu().numpy()
            if args.partition == "metis-renumbered":
                g, nb = fm.renumber_by_partition(g, epart, world)
                row_bounds = g.pdim * nb
            m = fm.localize(g, epart, rank)
            lab = [label]
            dist.broadcast_object_list(lab, 0)
            label = lab[0] + (", nodes renumbered by part, rows follow the parts" if row_bounds is not None else ", natural numbering, PETSC_DECIDE rows")
        else:
            m = fm.localize(g, mg.slab_partition(g, world), rank) if world > 1 else g
        del g
    u = mg.displacement_field(m)
    fx = api.Femx(device=lr)
    fx.set_mesh(m).build_plan()
    setup_s = time.time() - t0
    d_u = fx.device_buffer(u.nbytes).upload(u)
    d_f = fx.device_buffer(8 * fx.nrows)
    da = None
    if world > 1:
        da = par.DistributedAssembly(fx, dist, dev, peer=args.exchange == "peer", row_bounds=row_bounds if args.config == "c3" else None)
        vals_ptr = C.c_void_p(da.vals_local.data_ptr())
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    else:
        d_vals = fx.device_buffer(8 * fx.nnz)
        vals_ptr = d_vals.ptr

    def step():
        fx._ck(fx.L.femx_assemble_K_fint(fx.h, d_u.ptr, vals_ptr, d_f.ptr, api.DEVICE))
        ms = fx.last_kernel_ms
        if da is not None and not da.peer:
            ev0.record(); da.exchange(); ev1.record(); ev1.synchronize()
            ms += ev0.elapsed_time(ev1)
        return ms

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    w0 = time.perf_counter()
    dev_ms = sum(step() for _ in range(args.steps))
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - w0)
    stress_ms = 0.0
    if args.stress:   # nodal-averaged stress incl. the cross-rank sum (wall time: kernels + NCCL all_to_all of the shared nodes)
        d_s = fx.device_buffer(8 * m.cpdim * m.nlnds)
        ds = par.DistributedStress(fx, dist, dev) if world > 1 else None
        one = (lambda: ds.stress(d_u.ptr)) if ds else (lambda: fx._ck(fx.L.femx_stress(fx.h, d_u.ptr, d_s.ptr, api.ROW_MAJOR, api.DEVICE)))
        for _ in range(args.warmup):
            one()
        barrier()
        s0 = time.perf_counter()
        for _ in range(args.steps):
            one()
        barrier()
        stress_ms = 1e3 * (time.perf_counter() - s0) / args.steps
    t = torch.tensor([dev_ms, wall_ms, float(m.nels), float(fx.nnz), stress_ms], device=dev, dtype=torch.float64)
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    else:
        tmax = tsum = t
    if rank == 0:
        ms = float(tmax[0]) / args.steps
        print(json.dumps({"config": label, "n_gpus": world, "scaling": "strong", "elements": total, "ms_per_step": ms,
                          "wall_ms_per_step": float(tmax[1]) / args.steps, "elements_per_s": total / (ms * 1e-3),
                          "max_elements_per_rank": int(tmax[2]), "nnz_local_sum": int(tsum[3]),
                          "exchange": ("peer push kernel + neighbours' step barrier + receiver add (inside the timed call)" if da is not None and da.peer else "nccl all_to_all") if world > 1 else "none",
                          "recv_bytes_rank0": int(da.ep.recv_counts.sum()) * 8 if da is not None else 0,
                          "stress_wall_ms_per_call": float(tmax[4]) if args.stress else None,
                          "steps": args.steps, "warmup": args.warmup, "setup_s": round(setup_s, 1)}), flush=True)
    if da is not None:
        da.close()
    fx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
